// ref_driver_b.cpp -- TEST INFRASTRUCTURE (oracle/_ref).  Runs the reference's OWN,
// UNMODIFIED MIS kernel `drawTriangle` on the CPU: Sources/gpuRaytracer/shaders.metal
// (which includes Sources/CHeaders/shaderTypes.h) is #included from where it lies under
// /root/reference and compiled as C++ against oracle/metal_compat.  Nothing of it is
// copied.  The shader's literals stay: 6 camera rays, 300 MIS samples
// (shaders.metal:644,648).  Used to pin oracle/rt_oracle_modeb.c bit for bit.
//
// Built with -fpermissive: the shader declares a struct member `ray ray;`
// (shaders.metal:17), which C++ accepts only as a permissive extension.
#include "metal_compat/metal_stdlib"
#include "metal_compat/metal_raytracing"
#include "metal_compat/simd/simd.h"
#include <cstdint>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace ref_b {
#include "metal_compat/shader_scope.inc"
#include RT_REF_SHADERS_METAL
}

// One invocation per grid point, like dispatchThreads(width x height x 1)
// (Sources/gpuRaytracer/computeShader.swift:164-170).  pixels: RGBA8 (buffer 4),
// text: packed float3 per pixel (buffer 6); both full-image sized.
extern "C" int rt_ref_drawtriangle(const void *cameras, const void *materials, const void *lights,
                                   const void *vertices, uint32_t numTriangles,
                                   unsigned char *pixels, float *text, int width, int height, int row0, int rows) {
    using namespace metal;
    using namespace metal::raytracing;
    primitive_acceleration_structure accel;
    accel.vertices = (const float *)vertices;
    accel.triangleCount = numTriangles;
#pragma omp parallel for schedule(dynamic, 1)
    for (int y = row0; y < row0 + rows; y++)
        for (int x = 0; x < width; x++)
            ref_b::drawTriangle((const ref_b::CameraGPU *)cameras, (const ref_b::MaterialGPU *)materials,
                                (const ref_b::SquareLightGPU *)lights, (const float3 *)vertices, pixels,
                                accel, text, uint2((uint)x, (uint)y));
    return 0;
}

extern "C" void rt_ref_b_hash_random(uint32_t x, uint32_t y, uint32_t i, float out[2]) {
    metal::float2 r = ref_b::hashRandom(metal::uint2(x, y), i);
    out[0] = r.x; out[1] = r.y;
}
