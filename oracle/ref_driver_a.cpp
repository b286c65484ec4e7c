// ref_driver_a.cpp -- TEST INFRASTRUCTURE (oracle/_ref).  Runs the reference's OWN,
// UNMODIFIED kernel `pathTrace` on the CPU: RTrace/raytrace.metal (which imports
// RTrace/sampling.metal and RTrace/shaderTypes.h) is #included from where it lies under
// /root/reference and compiled as C++ against oracle/metal_compat.  Nothing of it is
// copied.  The shader's literals stay as they are: 400 samples, 3 bounces
// (raytrace.metal:24-25).  Used by tests/test_oracle_vs_ref.py to pin the C
// restatement (rt_oracle.c) bit for bit, and to generate tests/golden/.
#include "metal_compat/metal_stdlib"
#include "metal_compat/metal_raytracing"
#include "metal_compat/simd/simd.h"
#include <cstdint>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace ref_a {
#include "metal_compat/shader_scope.inc"
#include RT_REF_RAYTRACE_METAL
}

// One invocation of the kernel per grid point, exactly like
// dispatchThreads(width x height x 1) (RTrace/renderer.swift:137-141).
// out: float4 per pixel (the rgba16Float render target kept in fp32);
// offsets: the r32Uint randomTexture.  Rows [row0, row0+rows) are rendered; both
// arrays are full-image sized (width*height).
extern "C" int rt_ref_pathtrace(const void *cameras, const void *materials, const void *lights,
                                const void *vertices, uint32_t numTriangles, uint32_t *offsets,
                                float *out, int width, int height, int row0, int rows) {
    using namespace metal;
    using namespace metal::raytracing;
    texture2d<float, access::write> outputTexture;
    outputTexture.data = out; outputTexture.width = width; outputTexture.height = height; outputTexture.channels = 4;
    texture2d<unsigned int, access::read> randomTexture;
    randomTexture.data = offsets; randomTexture.width = width; randomTexture.height = height; randomTexture.channels = 1;
    primitive_acceleration_structure accel;
    accel.vertices = (const float *)vertices;
    accel.triangleCount = numTriangles;
    float debug[4] = { 0, 0, 0, 0 };
    // jobs of 32 grid points of one row, so that a few rows still keep every core busy
    const int chunks = (width + 31) / 32;
#pragma omp parallel for schedule(dynamic, 1)
    for (int job = 0; job < rows * chunks; job++)
        for (int y = row0 + job / chunks, x = (job % chunks) * 32, xe = x + 32 < width ? x + 32 : width; x < xe; x++)
            ref_a::pathTrace((const ref_a::CameraGPU *)cameras, (const ref_a::MaterialGPU *)materials,
                             (const ref_a::SquareLightGPU *)lights, (const float3 *)vertices,
                             outputTexture, randomTexture, accel, debug, uint3((uint)x, (uint)y, 0u));
    return 0;
}

// the reference's helper functions, for known-answer tests
extern "C" float rt_ref_halton(uint32_t i, uint32_t d) { return ref_a::halton(i, d); }
extern "C" uint32_t rt_ref_hash(uint32_t x) { return ref_a::hash(x); }
