/*
 * rt_oracle.h -- TEST INFRASTRUCTURE.  Public interface of liboracle.so, the CPU
 * restatement of the reference's per-pixel kernels.  Only tests/,
 * __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference) use it;
 * the product (gpuraytracer_b200/, include/) never links or loads it.
 *
 * Parity status: the reference has no tests, golden vectors or CPU code for this
 * path.  The restatement is pinned instead against
 * the reference's OWN shader source: oracle/_ref compiles the unmodified
 * RTrace/raytrace.metal + RTrace/sampling.metal as C++ (oracle/metal_compat) and
 * tests/test_oracle_vs_ref.py requires bit-identical images.  What stays
 * unpinned, and is defined in rt_detmath.h instead, is the closed-source part:
 * Metal's ray/triangle query and its fast-math library functions.
 */
#ifndef RT_ORACLE_H
#define RT_ORACLE_H

#include <stdint.h>
#include "../include/shaderTypes.h"
#include "../include/rtcuda.h"     /* RTLaunchParams only (plain struct) */

#ifdef __cplusplus
extern "C" {
#endif

/* Exact integer work counters (algorithmic FLOP formula). */
enum {
    RT_CNT_PATHS = 0,       /* camera samples = pixels * spp                  */
    RT_CNT_CLOSEST_RAYS,    /* closest-hit queries                            */
    RT_CNT_SHADOW_RAYS,     /* any-hit queries                                */
    RT_CNT_SHADED_BOUNCES,  /* non-emissive hits that were shaded             */
    RT_CNT_COUNT = 8
};

/* Mode A: kernel pathTrace, RTrace/raytrace.metal:11-111.
 * Buffers are the bytes the host uploads (shaderTypes.h layout).  `offsets` and
 * `out` hold rows [row0, row0+rows) only: element ((y-row0)*width + x).
 * out is float4 per pixel.  nthreads <= 0 means omp_get_max_threads().
 * Returns 0, or a negative RT_ERR_* for bad arguments. */
int rt_oracle_pathtrace(const CameraGPU *cameras, const MaterialGPU *materials,
                        const SquareLightGPU *lights, const simd_float3 *vertices,
                        const SphereGPU *spheres, const uint32_t *offsets,
                        float *out, const RTLaunchParams *params,
                        uint64_t counters[RT_CNT_COUNT], int nthreads);

/* Mode B: kernel drawTriangle, Sources/gpuRaytracer/shaders.metal:635-707.
 * pixels: RGBA8 (buffer 4), text: packed float3 (buffer 6), both tile-relative. */
int rt_oracle_drawtriangle(const CameraGPU *cameras, const MaterialGPU *materials,
                           const SquareLightGPU *lights, const simd_float3 *vertices,
                           uint8_t *pixels, float *text, const RTLaunchParams *params,
                           uint64_t counters[RT_CNT_COUNT], int nthreads);

/* Pieces exported for known-answer tests. */
float    rt_oracle_halton(uint32_t i, uint32_t d);            /* sampling.metal:107-122 */
/* Checker-only selectors in RTLaunchParams.variant beside RT_VARIANT_MIS (rt_oracle.c: "multiple importance sampling in
 * mode A"): the two strategies alone, each with weight one.  The library has no such forms. */
#define RT_ORACLE_MIS_LIGHT_ONLY  0x0800u
#define RT_ORACLE_MIS_COSINE_ONLY 0x1000u
/* per-light normals (float4 per light; NULL = the reference's (0,-1,0) literals): what the library takes as a buffer
 * bound at RT_INDEX_LIGHT_NORMALS.  A process-wide setting of the checker: set, trace, reset. */
void     rt_oracle_set_light_normals(const float *normals_float4_per_light);
uint32_t rt_oracle_hash(uint32_t x);                          /* sampling.metal:68-75   */
void     rt_oracle_camera_ray(const CameraGPU *cam, uint32_t x, uint32_t y, float jx, float jy,
                              float origin[3], float dir[3]); /* sampling.metal:125-157 */
void     rt_oracle_sample_area_light(const SquareLightGPU *light, float ux, float uy,
                                     const float pos[3], float color[3], float ldir[3],
                                     float *ldist);           /* sampling.metal:198-236 */
void     rt_oracle_sample_area_light_ex(const SquareLightGPU *light, float ux, float uy,
                                        const float pos[3], float color[3], float ldir[3],
                                        float *ldist, int light_extent);   /* with RT_VARIANT_LIGHT_EXTENT */
void     rt_oracle_cosine_hemisphere(float ux, float uy, const float normal[3], float out[3]);
                                                              /* sampling.metal:39-66   */
void     rt_oracle_sincos(float x, float *s, float *c);       /* rt_detmath.h           */
int      rt_oracle_intersect(const simd_float3 *vertices, uint32_t numTriangles,
                             const SphereGPU *spheres, uint32_t numSpheres,
                             const float o[3], const float d[3], float tmin, float tmax,
                             int acceptAny, float *distance, uint32_t *primitive_id);
int      rt_oracle_intersect_ordered(const simd_float3 *vertices, const uint32_t *ids, uint32_t n,
                                     const float o[3], const float d[3], float tmin, float tmax,
                                     float *distance, uint32_t *primitive_id);   /* bvh.cuh's order-independent rule */
/* image.swift:40-65: x2 exposure, Reinhard, gamma 2.2 -> RGBA8 */
void     rt_oracle_tonemap_rgba8(const float *rgba, uint8_t *out, uint64_t npixels);
void     rt_oracle_resolve_rgba16f(const float *rgba, uint16_t *out, uint64_t npixels, int divide_by_alpha);

/* ---- fixtures (restated input data) ---------------------------------------- */
/* Cornell box of RTrace/scene.swift:14-240: 36 triangles (108 vertices), their 36
 * materials, one light, one camera.  lightSize = 1.0 (RTrace) or 1.5 (mode B,
 * Sources/gpuRaytracer/main.swift:30-31).  Arrays must hold 108 / 36 / 1 / 1. */
void rt_fixture_cornell(int width, int height, float lightSize, CameraGPU *camera,
                        MaterialGPU *materials, SquareLightGPU *light, simd_float3 *vertices);
/* Config 4: room walls + 4 ceiling lights (each also 2 emissive triangles) +
 * numSpheres spheres on a jittered lattice.  Triangles: 10 + 8 = 18. */
void rt_fixture_spheres(int width, int height, uint32_t seed, uint32_t numSpheres,
                        CameraGPU *camera, MaterialGPU *materials /*18*/,
                        SquareLightGPU *lights /*4*/, simd_float3 *vertices /*54*/,
                        SphereGPU *spheres /*numSpheres*/);
/* Seeded stand-in for the arc4random() % 2^20 texture of renderer.swift:96-101:
 * offsets[(y-row0)*width+x] = hash(y*width+x + 0x9E3779B9*seed) & 0xFFFFF. */
void rt_fixture_offsets(uint32_t *offsets, int width, int row0, int rows, uint32_t seed);

/* ---- object groups of the sorted kernel (rt_groups.c) ------------------------ */
/* out: the 52 words of the prepared scene's "group" section (csrc/rt_scene_layout.h); returns K */
int  rt_oracle_groups(const simd_float3 *vertices, uint32_t T, const SquareLightGPU *lights, uint32_t L, uint32_t out[52]);
/* traces the frame and repeats every ray query against the quads group culling leaves: stats = queries,
 * mismatches (must be 0), triangle tests with / without culling, keysDual[16], keysSingle[16] */
int  rt_oracle_cullcheck(const CameraGPU *cameras, const MaterialGPU *materials, const SquareLightGPU *lights,
                         const simd_float3 *vertices, const uint32_t *offsets, float *out,
                         const RTLaunchParams *params, uint64_t stats[36], int nthreads);

/* mode B's form of the same check (every closest-hit query of a drawTriangle frame; the shell rule for the rays that are
 * segments towards a light sample point): stats = queries, mismatches (must be 0), triangle tests with / without culling,
 * segment queries, of those with both ends inside the shell box */
int  rt_oracle_cullcheck_modeb(const CameraGPU *cameras, const MaterialGPU *materials, const SquareLightGPU *lights,
                               const simd_float3 *vertices, uint8_t *pixels, float *text,
                               const RTLaunchParams *params, uint64_t stats[6], int nthreads);
/* the sphere grid's check (rt_sgrid.c): stats = queries, mismatches (must be 0), cells visited, sphere tests made, grid
 * usable; firstBad = the first failing query (o, d, tmax, any) */
int  rt_oracle_sgridcheck(const CameraGPU *cameras, const MaterialGPU *materials, const SquareLightGPU *lights,
                          const simd_float3 *vertices, const SphereGPU *spheres, const uint32_t *offsets, float *out,
                          const RTLaunchParams *params, uint64_t stats[5], float firstBad[8], int nthreads);

/* the group section of a scene WITH spheres (its triangles are grouped for the shell's sake only: word [2] stays 0), and the
 * check of the several-lights passes' shell rule (csrc/wave.cuh): stats = shadow queries, of those repeated without the
 * shell triangles, mismatches (must be 0) */
int  rt_oracle_groups_spheres(const simd_float3 *vertices, uint32_t T, const SquareLightGPU *lights, uint32_t L, uint32_t out[52]);
int  rt_oracle_shellcheck(const CameraGPU *cameras, const MaterialGPU *materials, const SquareLightGPU *lights,
                          const simd_float3 *vertices, const SphereGPU *spheres, const uint32_t *offsets, float *out,
                          const RTLaunchParams *params, uint64_t stats[3], int nthreads);

int  rt_oracle_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
