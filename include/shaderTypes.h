/*
 * shaderTypes.h -- host/device byte-layout contract for the path-tracing kernel.
 *
 * Replaces (drop-in, same type and field names, same sizes and offsets):
 *   reference RTrace/shaderTypes.h:13-94  ==  Sources/CHeaders/shaderTypes.h:13-94
 *
 * The reference header is Apple-only (it includes <simd/simd.h>).  This one
 * compiles with gcc, g++ and nvcc as well: when <simd/simd.h> is missing it
 * supplies simd_float3 / simd_float4 / simd_int2 with Apple's size and
 * alignment (16/16, 16/16, 8/8), so the buffers a Swift host uploads through
 * the Metal API and the buffers uploaded through include/rtcuda.h are the
 * same bytes.  Every size and offset of SURVEY.md section 8(a) row A9 is
 * checked at compile time at the bottom of this file.
 */
#ifndef ShaderTypes_h
#define ShaderTypes_h

#if defined(__has_include)
#  if __has_include(<simd/simd.h>)
#    include <simd/simd.h>
#    define RT_HAVE_APPLE_SIMD 1
#  endif
#endif

#ifndef RT_HAVE_APPLE_SIMD
/* Portable stand-ins.  simd_float3 is a 16-byte, 16-byte-aligned vector whose
 * fourth lane is padding, exactly like Apple's. */
#  if defined(__CUDACC__) || defined(__GNUC__)
#    define RT_ALIGNED(n) __attribute__((aligned(n)))
#  else
#    define RT_ALIGNED(n) __declspec(align(n))
#  endif
typedef struct RT_ALIGNED(16) { float x, y, z, _pad; } simd_float3;
typedef struct RT_ALIGNED(16) { float x, y, z, w; }    simd_float4;
typedef struct RT_ALIGNED(8)  { int x, y; }            simd_int2;
#endif

/* Surface description of one primitive; 48 bytes.
 * Filled by convertMaterial (reference RTrace/computeShader.swift:13-20). */
typedef struct {
    simd_float4 diffuse;      /* @0  rgb albedo, a unused                    */
    float       metallic;     /* @16 mode B only                             */
    float       roughness;    /* @20 mode B only                             */
    simd_float3 emissive;     /* @32 any component > 0 marks a light surface */
} MaterialGPU;

typedef struct {
    simd_float3 vertices[3];  /* @0  */
    MaterialGPU material;     /* @48 */
} TriangleGPU;                /* 96 bytes */

typedef struct {
    simd_float3 center;       /* @0  */
    MaterialGPU material;     /* @16 */
    float       radius;       /* @64 */
} SphereGPU;                  /* 80 bytes */

/* Pinhole camera; 64 bytes (reference convertCameras, computeShader.swift:242-253). */
typedef struct {
    simd_float3 position;      /* @0  */
    simd_float3 direction;     /* @16 */
    simd_float3 up;            /* @32 */
    simd_int2   resolution;    /* @48 */
    float       horizontalFov; /* @56 radians */
    float       ev100;         /* @60 */
} CameraGPU;

typedef struct {
    simd_float3 center;
    simd_float4 color;
    simd_float3 emittedRadiance;
    float       radius;        /* @48 */
} SphereLightGPU;              /* 64 bytes */

typedef struct {
    simd_float3 center;
    simd_float4 color;
    simd_float3 emittedRadiance;
    float       width;         /* @48 */
    float       height;        /* @52 */
    float       depth;         /* @56 */
} BoxLightGPU;                 /* 64 bytes */

/* Rectangular area light facing -y; 64 bytes
 * (reference convertSquareLight, computeShader.swift:33-41). */
typedef struct {
    simd_float3 center;          /* @0  */
    simd_float4 color;           /* @16 */
    simd_float3 emittedRadiance; /* @32 */
    float       width;           /* @48 */
    float       depth;           /* @52 */
} SquareLightGPU;

typedef struct {
    simd_float3 origin;
    simd_float3 direction;
    float       minDistance;   /* @32 */
    float       maxDistance;   /* @36 */
} RayGPU;                      /* 48 bytes */

typedef enum {
    Hit,
    HitLight,
    Miss
} IntersectionTypeGPU;

typedef struct {
    simd_float3 direction;
    float       pdf;           /* @16 */
    simd_float3 radiance;      /* @32 */
} SampleResultGPU;             /* 48 bytes */

typedef struct {
    simd_float3 tangent;
    simd_float3 bitangent;
} OrthonormalBasisGPU;         /* 32 bytes */

/* ---- layout checks ---------------------------------- */
#ifndef RT_SKIP_LAYOUT_CHECKS
#  include <stddef.h>
#  if defined(__cplusplus)
#    define RT_SA(c, m) static_assert(c, m)
#  else
#    define RT_SA(c, m) _Static_assert(c, m)
#  endif
RT_SA(sizeof(simd_float3) == 16 && sizeof(simd_float4) == 16 && sizeof(simd_int2) == 8, "simd sizes");
RT_SA(sizeof(MaterialGPU) == 48, "MaterialGPU size");
RT_SA(offsetof(MaterialGPU, metallic) == 16 && offsetof(MaterialGPU, roughness) == 20 &&
      offsetof(MaterialGPU, emissive) == 32, "MaterialGPU offsets");
RT_SA(sizeof(TriangleGPU) == 96 && offsetof(TriangleGPU, material) == 48, "TriangleGPU");
RT_SA(sizeof(SphereGPU) == 80 && offsetof(SphereGPU, material) == 16 &&
      offsetof(SphereGPU, radius) == 64, "SphereGPU");
RT_SA(sizeof(CameraGPU) == 64 && offsetof(CameraGPU, direction) == 16 && offsetof(CameraGPU, up) == 32 &&
      offsetof(CameraGPU, resolution) == 48 && offsetof(CameraGPU, horizontalFov) == 56 &&
      offsetof(CameraGPU, ev100) == 60, "CameraGPU");
RT_SA(sizeof(SphereLightGPU) == 64 && offsetof(SphereLightGPU, radius) == 48, "SphereLightGPU");
RT_SA(sizeof(BoxLightGPU) == 64 && offsetof(BoxLightGPU, width) == 48 && offsetof(BoxLightGPU, height) == 52 &&
      offsetof(BoxLightGPU, depth) == 56, "BoxLightGPU");
RT_SA(sizeof(SquareLightGPU) == 64 && offsetof(SquareLightGPU, color) == 16 &&
      offsetof(SquareLightGPU, emittedRadiance) == 32 && offsetof(SquareLightGPU, width) == 48 &&
      offsetof(SquareLightGPU, depth) == 52, "SquareLightGPU");
RT_SA(sizeof(RayGPU) == 48 && offsetof(RayGPU, minDistance) == 32 && offsetof(RayGPU, maxDistance) == 36, "RayGPU");
RT_SA(sizeof(SampleResultGPU) == 48 && offsetof(SampleResultGPU, pdf) == 16 &&
      offsetof(SampleResultGPU, radiance) == 32, "SampleResultGPU");
RT_SA(sizeof(OrthonormalBasisGPU) == 32, "OrthonormalBasisGPU");
#endif

#endif /* ShaderTypes_h */
