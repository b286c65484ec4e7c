//
//  image.swift -- saveTextureToImage (RTrace/image.swift:15-100) over librtcuda.so.
//
//  UNVERIFIED: no Swift toolchain exists in the image this was written in; never compiled.  The C++ twin is
//  gpuraytracer_b200/csrc/host_renderer.cpp (saveTextureToImage), tested on a B200 against the oracle's tone-mapped image.
//
//  What changed against the reference (R = /root/reference/RTrace/image.swift):
//    R:25-38  texture.getBytes into [Float16]            -> the tone map runs on the device: rtResolveToRGBA8 does, per channel,
//    R:40-65  the exposure / Reinhard / gamma loop          Float(Float16(v)) * 2 -> v / (v + 1) -> pow(v, 1 / 2.2) -> clamp -> UInt8(v * 255),
//                                                           i.e. R:48-59 on what the rgba16Float target would have held
//    R:67-99  CGImage + CGImageDestination                -> unchanged: savePixelArrayToImage (R:102-157) is called with the bytes
//  A host that wants to keep R:25-65 exactly as written calls rtResolveToRGBA16F instead and reads the half4 image with
//  rtBufferContents: those are the bytes texture.getBytes returned (8 per pixel, round-to-nearest-even).
//
import Foundation
import CShaderTypes

func saveTextureToImage(device: OpaquePointer?, queue: OpaquePointer?, texture: OpaquePointer?, width: Int, height: Int,
                        fileName: String) throws {
    // R:16-23: the reference reads width / height / pixelFormat off the MTLTexture; the library's texture is a linear
    // buffer (rtTextureBuffer), so the caller passes the size it created the texture with
    guard let storage = rtTextureBuffer(texture), rtBufferLength(storage) >= width * height * 16 else {
        throw RTError(call: "rtTextureBuffer", code: -1, text: "render target smaller than \(width) x \(height) float4")
    }
    let pixelCount = width * height
    var rgba: OpaquePointer?                                             // RTBuffer*, 4 bytes per pixel
    try rt(rtBufferCreate(device, nil, pixelCount * 4, &rgba), "rtBufferCreate(rgba8)")
    defer { rtBufferRelease(rgba) }
    try rt(rtResolveToRGBA8(queue, storage, 0, rgba, 0, UInt64(pixelCount), 0), "rtResolveToRGBA8")   // R:40-65 on the GPU
    guard let bytes = rtBufferContents(rgba) else {                      // waits for the queue, then maps the result (R:35 getBytes)
        throw RTError(call: "rtBufferContents", code: -6, text: String(cString: rtLastErrorString()))
    }
    let rgbaData = Array(UnsafeBufferPointer(start: bytes.bindMemory(to: UInt8.self, capacity: pixelCount * 4), count: pixelCount * 4))
    savePixelArrayToImage(pixels: rgbaData, width: width, height: height, fileName: fileName)      // R:102-157, unchanged
}
