//
//  computeShader.swift -- the mode-B host driver `drawTriangle` (Sources/gpuRaytracer/computeShader.swift:99-189)
//  and setupAccelerationStructures (:45-97) over librtcuda.so.
//
//  UNVERIFIED: no Swift toolchain exists in the image this was written in; never compiled.  The same call sequence is
//  tested on a B200 as Python (gpuraytracer_b200/compute_shader.py: draw_triangle; tests/test_gpu_modeb.py compares its
//  float3 buffer bit for bit with shaders.metal compiled as C++).
//
//  Kept from the reference file as they are (not repeated here): convertMaterial :13-20, convertTriangle :22-31,
//  convertSquareLight :33-41, writeDebugArrayToFile :191-210, convertSpheres :212-220, convertCameras :222-233,
//  generateRandomArray :235-.  Replaced (R = the reference file):
//    R:45-97    setupAccelerationStructures: descriptor + scratch buffer + encoder.build -> rtAccelerationStructureCreate
//               (the library builds its structure -- object groups, or a BVH above 1024 triangles -- itself)
//    R:112-123  makeLibrary(URL:) + makeFunction(name: "drawTriangle")              -> rtPipelineCreate(device, "drawTriangle")
//    R:125-136  intersection function table, makeComputePipelineState, command queue/buffer -> rtQueueCreate, rtCommandBegin
//    R:138-149  makeBuffer x 6                                                       -> rtBufferCreate x 6
//    R:151-159  setBuffer 0,1,2,3,4 / setAccelerationStructure 5 / setBuffer 6       -> rtSetBuffer / rtSetAccelerationStructure, same indices
//    R:162-174  dispatchThreads(w x h, 8 x 8) / commit / waitUntilCompleted          -> rtDispatchThreads / rtCommit / rtWaitUntilCompleted
//    R:176-188  debugBuffer.contents() / pixelsBuffer.contents()                     -> rtBufferContents
//  The `device:` parameter becomes the RTDevice* of rtDeviceCreate; the MTLAccelerationStructure the RTAccel*.
//
import Foundation
import CShaderTypes

struct RTError: Error { let call: String; let code: Int32; let text: String }
@inline(__always) func rt(_ code: Int32, _ call: String) throws {
    if code != 0 /* RT_OK */ { throw RTError(call: call, code: code, text: String(cString: rtLastErrorString())) }
}

/// R:45-97.  Returns the RTAccel* and the vertex buffer it records (the reference makes a private vertex buffer too, :52-58).
func setupAccelerationStructures(device: OpaquePointer?, triangles: [Triangle]) throws -> (accel: OpaquePointer?, vertices: OpaquePointer?) {
    var allVertices: [simd_float3] = []                                                       // R:48-51
    for triangle in triangles { allVertices.append(contentsOf: triangle.vertices) }
    var vertexBuffer: OpaquePointer?
    try rt(rtBufferCreate(device, allVertices, allVertices.count * MemoryLayout<simd_float3>.size, &vertexBuffer), "rtBufferCreate(vertices)")
    var accel: OpaquePointer?
    try rt(rtAccelerationStructureCreate(device, vertexBuffer, 0, MemoryLayout<simd_float3>.stride,   // R:62-67
                                         UInt32(triangles.count), &accel), "rtAccelerationStructureCreate")
    return (accel, vertexBuffer)
}

func drawTriangle(device: OpaquePointer?, cameras: [Camera], triangles: [Triangle], squareLights: [SquareLight], pixels: [UInt8],
                  accelerationStructure: OpaquePointer?) throws -> [UInt8] {
    let camerasGPU = convertCameras(cameras: cameras)                                        // R:101-110, unchanged
    let materialsGPU = triangles.map { convertMaterial(material: $0.material) }
    let squareLightsGPU = [convertSquareLight(squareLight: squareLights[0])]
    var allVertices: [simd_float3] = []
    for triangle in triangles { allVertices.append(contentsOf: triangle.vertices) }

    var pipeline: OpaquePointer?, queue: OpaquePointer?, command: OpaquePointer?
    try rt(rtPipelineCreate(device, "drawTriangle", &pipeline), "rtPipelineCreate")          // R:112-134
    defer { rtPipelineRelease(pipeline) }
    try rt(rtQueueCreate(device, &queue), "rtQueueCreate")                                   // R:135
    defer { rtQueueRelease(queue) }

    var cameraBuffer: OpaquePointer?, materialsBuffer: OpaquePointer?, pixelsBuffer: OpaquePointer?
    var squareLightsBuffer: OpaquePointer?, vertexBuffer: OpaquePointer?, debugBuffer: OpaquePointer?
    try rt(rtBufferCreate(device, camerasGPU, camerasGPU.count * MemoryLayout<CameraGPU>.size, &cameraBuffer), "camera")           // R:138-149
    try rt(rtBufferCreate(device, materialsGPU, materialsGPU.count * MemoryLayout<MaterialGPU>.size, &materialsBuffer), "materials")
    try rt(rtBufferCreate(device, nil, pixels.count * MemoryLayout<UInt8>.size, &pixelsBuffer), "pixels")
    try rt(rtBufferCreate(device, squareLightsGPU, squareLightsGPU.count * MemoryLayout<SquareLightGPU>.size, &squareLightsBuffer), "lights")
    try rt(rtBufferCreate(device, allVertices, allVertices.count * MemoryLayout<simd_float3>.size, &vertexBuffer), "vertices")
    try rt(rtBufferCreate(device, nil, pixels.count * MemoryLayout<simd_float3>.size, &debugBuffer), "debug")     // (the reference over-allocates: 16 B per BYTE of pixels)
    defer { for b in [cameraBuffer, materialsBuffer, pixelsBuffer, squareLightsBuffer, vertexBuffer, debugBuffer] { rtBufferRelease(b) } }

    try rt(rtCommandBegin(queue, pipeline, &command), "rtCommandBegin")                      // R:136, 151-152
    defer { rtCommandRelease(command) }
    try rt(rtSetBuffer(command, cameraBuffer, 0, 0), "rtSetBuffer(0)")                       // R:153-159
    try rt(rtSetBuffer(command, materialsBuffer, 0, 1), "rtSetBuffer(1)")
    try rt(rtSetBuffer(command, squareLightsBuffer, 0, 2), "rtSetBuffer(2)")
    try rt(rtSetBuffer(command, vertexBuffer, 0, 3), "rtSetBuffer(3)")
    try rt(rtSetBuffer(command, pixelsBuffer, 0, 4), "rtSetBuffer(4)")
    try rt(rtSetAccelerationStructure(command, accelerationStructure, 5), "rtSetAccelerationStructure(5)")
    try rt(rtSetBuffer(command, debugBuffer, 0, 6), "rtSetBuffer(6)")

    let width = Int32(camerasGPU[0].resolution.x), height = Int32(camerasGPU[0].resolution.y)    // R:162-164
    try rt(rtDispatchThreads(command, width, height, 1, 8, 8, 1), "rtDispatchThreads")       // R:168-170
    try rt(rtCommit(command), "rtCommit")                                                     // R:171-172
    try rt(rtWaitUntilCompleted(command), "rtWaitUntilCompleted")                            // R:173-174

    // R:176-182: the kernel packs float3 tightly (12 bytes per pixel, shaders.metal:627-633)
    let debugCount = Int(width) * Int(height)
    if let debugPtr = rtBufferContents(debugBuffer) {
        let f = debugPtr.bindMemory(to: Float.self, capacity: debugCount * 3)
        let debugArray = (0..<debugCount).map { simd_float3(f[3 * $0], f[3 * $0 + 1], f[3 * $0 + 2]) }
        writeDebugArrayToFile(debugArray: debugArray, width: Int(width), height: Int(height), fileName: "debugOutput.txt")
    }
    // R:184-188
    guard let pixelPtr = rtBufferContents(pixelsBuffer) else {
        throw RTError(call: "rtBufferContents", code: -6, text: String(cString: rtLastErrorString()))
    }
    return Array(UnsafeBufferPointer(start: pixelPtr.bindMemory(to: UInt8.self, capacity: pixels.count), count: pixels.count))
}
