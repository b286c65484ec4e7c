// simd/simd.h -- TEST INFRASTRUCTURE (oracle/_ref).  Inside a Metal translation unit
// <simd/simd.h> makes simd_float3 etc. the shader's own vector types; the reference's
// RTrace/shaderTypes.h (included unmodified by the shaders) needs exactly these names.
#ifndef RT_METAL_COMPAT_SIMD
#define RT_METAL_COMPAT_SIMD
#include "../metal_stdlib"
typedef metal::float3 simd_float3;
typedef metal::float4 simd_float4;
typedef metal::int2   simd_int2;
#endif
