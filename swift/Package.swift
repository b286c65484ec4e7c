// swift-tools-version:5.7
// UNVERIFIED: no Swift toolchain exists in the image this was written in (swift / swiftc absent); never compiled.
//
// The reference's Package.swift (Package.swift:16-27) with ONE change: the C module CShaderTypes now also carries
// rtcuda.h -- the C ABI of librtcuda.so, which stands where the Metal objects stood -- and links the library.
// Sources/CHeaders/{shaderTypes.h, rtcuda.h} are links to this repo's include/ (shaderTypes.h is byte-compatible
// with the reference's: same names, sizes, offsets).  Everything the reference's host compiles besides the files
// in this directory tree stays as it is: scene.swift, main.swift, the converters of computeShader.swift.
import PackageDescription

let package = Package(
    name: "tempShader",
    platforms: [ .macOS(.v13) ],
    products: [ .executable(name: "gpuRaytracer", targets: ["gpuRaytracer"]) ],
    targets: [
        .target(
            name: "CShaderTypes",
            path: "Sources/CHeaders",
            sources: ["shaderTypes.h", "rtcuda.h"],
            publicHeadersPath: ".",
            linkerSettings: [
                .linkedLibrary("rtcuda"),                                   // gpuraytracer_b200/librtcuda.so
                .unsafeFlags(["-L", "../gpuraytracer_b200", "-Xlinker", "-rpath", "-Xlinker", "../gpuraytracer_b200"])
            ]
        ),
        .executableTarget(
            name: "gpuRaytracer",
            dependencies: ["CShaderTypes"],
            path: "Sources/gpuRaytracer"
            // (the reference excludes shaders.metal / shaders.air / MyMetalLib.metallib here: the kernels now live
            //  inside librtcuda.so as an embedded sm_100a cubin, there is nothing to exclude or to load at run time)
        )
    ]
)
