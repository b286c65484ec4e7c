//
//  renderer.swift -- the reference's `Renderer` (RTrace/renderer.swift:9-146) over librtcuda.so.
//
//  UNVERIFIED: no Swift toolchain exists in the image this was written in; never compiled.  The same sequence of
//  rt* calls IS compiled and tested on a B200 as C++ (gpuraytracer_b200/csrc/host_renderer.cpp,
//  tests/test_gpu_parity.py::test_rtrace_executable_writes_the_oracles_image) and as Python (gpuraytracer_b200/renderer.py).
//
//  What changed against the reference file, line by line (R = /root/reference/RTrace/renderer.swift):
//    R:7       import Metal                              -> import CShaderTypes (shaderTypes.h + rtcuda.h)
//    R:10-25   MTLDevice / MTLComputePipelineState / MTLCommandQueue / MTLBuffer / MTLTexture /
//              MTLAccelerationStructure members          -> opaque RT* handles (OpaquePointer?)
//    R:30      MTLCreateSystemDefaultDevice()!           -> rtDeviceCreate
//    R:32-43   Bundle url + makeLibrary + makeFunction + makeComputePipelineState
//                                                        -> rtPipelineCreate(device, "pathTrace") (the kernels are embedded)
//    R:44      makeCommandQueue()                        -> rtQueueCreate
//    R:61-72   makeBuffer(bytes:length:options:)         -> rtBufferCreate (copies the bytes, like storageModeShared)
//    R:74-93   two MTLTextureDescriptors + makeTexture   -> rtTextureCreate (RGBA32_FLOAT stands for rgba16Float; R32_UINT)
//    R:103-110 randomTexture.replace(region:...)         -> rtTextureReplaceRegion
//    R:113     setupAccelerationStructures(...)          -> rtAccelerationStructureCreate
//    R:118-144 command buffer / encoder / set* / dispatchThreads / commit / waitUntilCompleted
//                                                        -> rtCommandBegin / rtSet* / rtDispatchThreads / rtCommit / rtWaitUntilCompleted
//    R:146     saveTextureToImage(texture:fileName:)     -> the same call, image.swift of this directory
//  Unchanged and still used from the reference: initCornellBox (scene.swift), convertCameras / convertMaterial /
//  convertSquareLight (computeShader.swift:13-41,242-253), Camera, outputFileName (main.swift).
//  Failure: the reference force-unwraps and fatalError()s; every rt* call returns a code, turned into a throw here.
//
import Foundation
import CShaderTypes

public struct RTError: Error, CustomStringConvertible {
    public let call: String
    public let code: Int32
    public let text: String
    public var description: String { "\(call) failed (\(code)): \(text)" }
}

@inline(__always) func rt(_ code: Int32, _ call: String) throws {
    if code != 0 /* RT_OK */ { throw RTError(call: call, code: code, text: String(cString: rtLastErrorString())) }
}

public class Renderer {
    var device: OpaquePointer?                 // RTDevice*    (R:10 MTLDevice)
    var computePipeline: OpaquePointer?        // RTPipeline*  (R:12)
    var commandQueue: OpaquePointer?           // RTQueue*     (R:13)

    var cameraBuffer: OpaquePointer?           // RTBuffer*    (R:15-20)
    var materialsBuffer: OpaquePointer?
    var lightsBuffer: OpaquePointer?
    var vertexBuffer: OpaquePointer?
    var debugBuffer: OpaquePointer?

    var randomTexture: OpaquePointer?          // RTTexture*   (R:22-23)
    var renderTexture: OpaquePointer?

    var accelerationStructure: OpaquePointer?  // RTAccel*     (R:25)

    let camera: Camera
    let triangleCount: UInt32

    /// BASELINE configs change what the shader hard-codes (raytrace.metal:24-25); nil = the reference's 400 spp / 3 bounces.
    public var samplesPerPixel: UInt32? = nil
    public var bounces: Int32? = nil

    init() throws {
        try rt(rtDeviceCreate(0, &device), "rtDeviceCreate")                                      // R:30
        try rt(rtPipelineCreate(device, "pathTrace", &computePipeline), "rtPipelineCreate")      // R:32-43
        try rt(rtQueueCreate(device, &commandQueue), "rtQueueCreate")                            // R:44

        let cornellBoxScene = initCornellBox()                                                    // R:46
        let camerasGPU = convertCameras(cameras: [cornellBoxScene.camera])                        // R:49
        let materialsGPU = cornellBoxScene.triangles.map { convertMaterial(material: $0.material) }
        let squareLightsGPU = [convertSquareLight(squareLight: cornellBoxScene.light)]
        var allVertices: [simd_float3] = []                                                       // R:54-57
        for triangle in cornellBoxScene.triangles { allVertices.append(contentsOf: triangle.vertices) }
        camera = cornellBoxScene.camera
        triangleCount = UInt32(cornellBoxScene.triangles.count)

        // R:61-72  (the library copies the bytes: the arrays may go away)
        try rt(rtBufferCreate(device, camerasGPU, camerasGPU.count * MemoryLayout<CameraGPU>.size, &cameraBuffer), "rtBufferCreate(camera)")
        try rt(rtBufferCreate(device, materialsGPU, materialsGPU.count * MemoryLayout<MaterialGPU>.size, &materialsBuffer), "rtBufferCreate(materials)")
        try rt(rtBufferCreate(device, squareLightsGPU, squareLightsGPU.count * MemoryLayout<SquareLightGPU>.size, &lightsBuffer), "rtBufferCreate(lights)")
        try rt(rtBufferCreate(device, allVertices, allVertices.count * MemoryLayout<simd_float3>.size, &vertexBuffer), "rtBufferCreate(vertices)")
        try rt(rtBufferCreate(device, nil, camera.pixelCount() * MemoryLayout<simd_float3>.size, &debugBuffer), "rtBufferCreate(debug)")

        let width = Int32(camera.resolution.x), height = Int32(camera.resolution.y)
        // R:74-82: the render target.  fp32 here (SURVEY 8g.3); rtResolveToRGBA16F gives the reference's rgba16Float bytes
        try rt(rtTextureCreate(device, 1 /* RT_FORMAT_RGBA32_FLOAT */, width, height, &renderTexture), "rtTextureCreate(render)")
        // R:84-92
        try rt(rtTextureCreate(device, 2 /* RT_FORMAT_R32_UINT */, width, height, &randomTexture), "rtTextureCreate(random)")

        var randomData = [UInt32](repeating: 0, count: Int(width) * Int(height))                  // R:94-101
        for i in 0..<randomData.count { randomData[i] = UInt32(arc4random() % (1024 * 1024)) }
        try rt(rtTextureReplaceRegion(randomTexture, 0, 0, width, height, randomData,             // R:103-110
                                      Int(width) * MemoryLayout<UInt32>.size), "rtTextureReplaceRegion")

        // R:113  setupAccelerationStructures(device:triangles:); stride = MemoryLayout<simd_float3>.stride = 16 (computeShader.swift:66)
        try rt(rtAccelerationStructureCreate(device, vertexBuffer, 0, MemoryLayout<simd_float3>.stride,
                                             triangleCount, &accelerationStructure), "rtAccelerationStructureCreate")
    }

    func draw() throws {
        var command: OpaquePointer?                                                               // RTCommand*: R:118-121
        try rt(rtCommandBegin(commandQueue, computePipeline, &command), "rtCommandBegin")
        defer { rtCommandRelease(command) }
        try rt(rtSetBuffer(command, cameraBuffer, 0, 0), "rtSetBuffer(0)")                       // R:122
        try rt(rtSetBuffer(command, materialsBuffer, 0, 1), "rtSetBuffer(1)")                    // R:123
        try rt(rtSetBuffer(command, lightsBuffer, 0, 2), "rtSetBuffer(2)")                       // R:124
        try rt(rtSetBuffer(command, vertexBuffer, 0, 3), "rtSetBuffer(3)")                       // R:125
        try rt(rtSetAccelerationStructure(command, accelerationStructure, 5), "rtSetAccelerationStructure(5)")  // R:127
        try rt(rtSetBuffer(command, debugBuffer, 0, 6), "rtSetBuffer(6)")                        // R:128
        try rt(rtSetTexture(command, renderTexture, 0), "rtSetTexture(0)")                       // R:129
        try rt(rtSetTexture(command, randomTexture, 1), "rtSetTexture(1)")                       // R:130

        let width = Int32(camera.resolution.x), height = Int32(camera.resolution.y)              // R:133-134
        if samplesPerPixel != nil || bounces != nil {
            var params = RTLaunchParams()
            try rt(rtDefaultParams(width, height, &params), "rtDefaultParams")                   // 400 spp, 3 bounces
            if let spp = samplesPerPixel { params.spp = spp }
            if let b = bounces { params.bounces = b }
            params.numTriangles = triangleCount
            try rt(rtSetParams(command, &params), "rtSetParams")
        }
        try rt(rtDispatchThreads(command, width, height, 1, 8, 8, 1), "rtDispatchThreads")       // R:137-141
        try rt(rtCommit(command), "rtCommit")                                                     // R:142-143
        try rt(rtWaitUntilCompleted(command), "rtWaitUntilCompleted")                            // R:144

        try saveTextureToImage(device: device, queue: commandQueue, texture: renderTexture, width: Int(width), height: Int(height),
                               fileName: outputFileName)                                          // R:146
    }

    deinit {   // ARC released the Metal objects; children retain their parents in the library, any order is safe
        rtAccelerationStructureRelease(accelerationStructure)
        rtTextureRelease(renderTexture); rtTextureRelease(randomTexture)
        for b in [cameraBuffer, materialsBuffer, lightsBuffer, vertexBuffer, debugBuffer] { rtBufferRelease(b) }
        rtQueueRelease(commandQueue); rtPipelineRelease(computePipeline); rtDeviceRelease(device)
    }
}
