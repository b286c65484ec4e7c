/*
 * rtcuda.h -- C ABI of librtcuda.so: the B200 (sm_100a) stand-in for the Metal
 * objects that drive the reference's path-tracing kernel.
 *
 * The reference host (Swift) talks to the GPU through MTLDevice, MTLLibrary,
 * MTLComputePipelineState, MTLCommandQueue, MTLCommandBuffer,
 * MTLComputeCommandEncoder, MTLBuffer, MTLTexture and MTLAccelerationStructure
 * (RTrace/renderer.swift:29-146, RTrace/computeShader.swift:45-97,
 * Sources/gpuRaytracer/computeShader.swift:99-189).  Each entry point below
 * replaces one of those calls and cites it; INTEGRATION.md shows the Swift side.
 * Plain C: opaque handles, pointers and sizes only.  Buffer contents use the
 * byte layout of include/shaderTypes.h.
 *
 * Conventions
 *   - every function returns int: 0 = RT_OK, > 0 = a CUresult from the CUDA
 *     driver, < 0 = one of the RT_ERR_* codes below.  Nothing aborts or throws
 *     across this boundary (the reference force-unwraps and fatalError()s,
 *     renderer.swift:30-44); rtLastErrorString() gives the text of the last
 *     failure on the calling thread.
 *   - handles are owned by the caller and freed with the matching rt*Release.
 *     They are not thread-safe; different RTDevice objects may be driven from
 *     different host threads.
 *   - there is NO CPU fallback: without a CUDA driver and an sm_100 GPU,
 *     rtDeviceCreate fails with RT_ERR_NO_DRIVER / a CUresult.
 */
#ifndef RTCUDA_H
#define RTCUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_ABI_VERSION 1

enum {
    RT_OK                   =  0,
    RT_ERR_INVALID_ARGUMENT = -1,   /* NULL handle, bad size, bad index          */
    RT_ERR_UNKNOWN_KERNEL   = -2,   /* makeFunction(name:) would have returned nil */
    RT_ERR_NO_DRIVER        = -3,   /* libcuda.so.1 cannot be loaded              */
    RT_ERR_MISSING_BINDING  = -4,   /* a buffer the kernel reads was never bound  */
    RT_ERR_OUT_OF_RANGE     = -5,   /* binding too small for the launch           */
    RT_ERR_STATE            = -6,   /* call order (e.g. wait before commit)       */
    RT_ERR_UNSUPPORTED      = -7,   /* e.g. bounces > 4 (primes[24] would overrun) */
    RT_ERR_OUT_OF_MEMORY    = -8,
    RT_ERR_DEVICE_TIMEOUT   = -9    /* a kernel gave up a bounded wait (chunk hand-over of the ticketed forms): the image is invalid */
};

typedef struct RTDevice   RTDevice;    /* MTLDevice                                   */
typedef struct RTPipeline RTPipeline;  /* MTLLibrary + MTLFunction + MTLComputePipelineState */
typedef struct RTQueue    RTQueue;     /* MTLCommandQueue (one CUDA stream)           */
typedef struct RTCommand  RTCommand;   /* MTLCommandBuffer + MTLComputeCommandEncoder */
typedef struct RTBuffer   RTBuffer;    /* MTLBuffer, storageModeShared semantics      */
typedef struct RTTexture  RTTexture;   /* MTLTexture (stored as a linear buffer)      */
typedef struct RTAccel    RTAccel;     /* MTLAccelerationStructure (geometry record)  */
typedef struct RTGroup    RTGroup;     /* new: several devices rendering row tiles    */

/* Argument-table slots.  0..6 are the reference's own [[buffer(n)]] indices
 * (raytrace.metal:11-18, shaders.metal:635-637); textures 0 and 1 of pathTrace
 * are bound with rtSetTexture and land on RT_INDEX_OUTPUT / RT_INDEX_OFFSETS. */
enum {
    RT_INDEX_CAMERA    = 0,    /* CameraGPU[1]                                  */
    RT_INDEX_MATERIALS = 1,    /* MaterialGPU[T]                                */
    RT_INDEX_LIGHTS    = 2,    /* SquareLightGPU[L]                             */
    RT_INDEX_VERTICES  = 3,    /* simd_float3[3T]                               */
    RT_INDEX_PIXELS    = 4,    /* drawTriangle only: RGBA8                      */
    RT_INDEX_ACCEL     = 5,    /* acceleration structure                        */
    RT_INDEX_DEBUG     = 6,    /* pathTrace: bound, never written; drawTriangle: float3 text buffer */
    RT_INDEX_SPHERES   = 7,    /* extension (config 4): SphereGPU[S]            */
    RT_INDEX_LIGHT_NORMALS = 8, /* extension, optional: float4[L], xyz = the light's normal (normalised by
                                  the library); unbound = the reference's literal (0,-1,0), sampling.metal:205.
                                  A side table, so that SquareLightGPU keeps the reference's layout */
    RT_INDEX_OUTPUT    = 16,   /* texture(0) of pathTrace: float4 per pixel     */
    RT_INDEX_OFFSETS   = 17,   /* texture(1) of pathTrace: uint32 per pixel     */
    RT_INDEX_COUNT     = 18
};

enum { RT_FORMAT_RGBA32_FLOAT = 1, RT_FORMAT_R32_UINT = 2 };
/* Kernel variants of pathTrace: schedules of the one reference kernel, all giving bit-identical images.
 *   DEFAULT             the library's choice, a rule table (csrc/rt_plan.cpp; tests/test_plan_cpu.py pins it per BASELINE
 *                       config): SORTED_BLOCK for scenes that can be cut into object groups (quads, <= 64 triangles, no
 *                       spheres, one light: the Cornell scene), its ticketed form when a launch has fewer than 8 resident
 *                       waves of blocks and >= 64 spp (one rank's share of a multi-GPU frame), its wide form below 6 spp
 *                       (progressive rendering); WAVE / TILES / WIDE likewise for every other scene; BVH above 1024
 *                       triangles; SMEM when the path queues do not fit beside the scene
 *   SMEM                one thread walks one pixel's samples (the independent second implementation the tests compare with)
 *   CONST               the same with the primitives in the constant bank
 *   WAVE                warp-level path compaction (csrc/wave.cuh)
 *   TILES               WAVE with one block per (tile, sample chunk) unit, named by a ticket from a counter
 *   WIDE                WAVE for launches with fewer than 6 samples per pixel: a warp takes 6 / spp tiles side by side
 *   SORTED              WAVE with the paths of a round sorted by the object groups their rays can reach, so that full warps
 *                       test only those groups' quads (csrc/sorted.cuh; per-warp sort)
 *   SORTED_BLOCK        the sort taken over the block's four warps (csrc/sortedblk.cuh)
 *   SORTED_TILES, SORTED_BLOCK_TILES   their ticketed forms
 *   BVH                 one thread per pixel, every ray query through a bounding-volume hierarchy in global memory
 *                       (csrc/bvh.cuh), built on the device when the acceleration structure is created
 * A form that does not apply to the launch (scene cannot be grouped, too few samples, counting requested) falls back as
 * the table says.  Flags above the low byte: COUNT_RAYS (below); LIGHT_EXTENT: sampleAreaLight takes its sample over the
 * light's own width x depth, right = (width/2,0,0), up = (0,0,depth/2), instead of the 0.5 x 0.5 the reference hard-codes
 * (sampling.metal:207-209, the TODO SURVEY N3 names); off = the reference's literals.  Every kernel form honours it, the
 * image changes with it (the oracle takes the same flag).
 * MIS (the reference's own TODO, the commented-out block of sampling.metal:238-260 that mode B carries out):
 * the direct light of every shaded bounce is estimated by the light sample AND by the cosine-sampled continuation ray,
 * combined with the balance heuristic (powerHeuristic, beta = 1, over calculateSquareLightPdf and
 * calculateCosineWeightedPdf: shaders.metal:132-137, 315-326, 376-380).  Same expectation as the flag-off light term,
 * lower variance near the light; the continuation ray is then traced after the last bounce too, and an emissive hit
 * after a shaded bounce ends the path without raytrace.metal:59's `accumulated = emissive` (oracle/rt_oracle.c states
 * the estimator).  Forms: SMEM, WAVE, SORTED_BLOCK, BVH; the others fall back to those.  Not with COUNT_RAYS. */
enum { RT_VARIANT_DEFAULT = 0, RT_VARIANT_SMEM = 1, RT_VARIANT_CONST = 2, RT_VARIANT_WAVE = 3, RT_VARIANT_TILES = 16, RT_VARIANT_WIDE = 17, RT_VARIANT_SORTED = 18, RT_VARIANT_SORTED_TILES = 19, RT_VARIANT_SORTED_BLOCK = 20, RT_VARIANT_SORTED_BLOCK_TILES = 21, RT_VARIANT_BVH = 22,
       RT_VARIANT_COUNT_RAYS = 0x100, RT_VARIANT_LIGHT_EXTENT = 0x200, RT_VARIANT_MIS = 0x400 };

/* What the reference hard-codes in the shader (raytrace.metal:24-25,
 * shaders.metal:644,648) plus the tile / progressive controls BASELINE.json's
 * configs need.  rtSetParams is optional: without it a command uses
 * rtDefaultParams(): the reference's literals over the whole dispatch grid. */
typedef struct RTLaunchParams {
    int32_t  width, height;    /* image size = dispatch grid of the reference        */
    int32_t  row0, rows;       /* this launch renders global rows [row0, row0+rows);
                                  the output / offsets / pixels bindings start at row0 */
    uint32_t spp;              /* pathTrace: samples per pixel (reference: 400)      */
    uint32_t firstSample;      /* sample index n runs firstSample..firstSample+spp-1 */
    int32_t  bounces;          /* pathTrace: 3 in the reference; 1..4                */
    int32_t  accumulate;       /* 0: out = mean over spp, alpha 1 (raytrace.metal:106-109)
                                  1: out.xyz += sum of the spp samples (continuing the
                                     running fp32 sum stored in out), out.w = firstSample+spp */
    uint32_t numTriangles;     /* T                                                  */
    uint32_t numSpheres;       /* S (0 in the reference scenes)                      */
    uint32_t numLights;        /* L (reference kernels read lights[0] only)          */
    uint32_t cameraRays;       /* drawTriangle: camera rays per pixel (reference: 6) */
    uint32_t misSamples;       /* drawTriangle: MIS samples, split in 3 (reference: 300) */
    uint32_t variant;          /* low byte: 0 = tuned default, else a kernel variant for profiling
                                  (DESIGN.md).  RT_VARIANT_COUNT_RAYS (0x100): also add the exact
                                  work counters {paths, closest rays, shadow rays, shaded bounces}
                                  as 4 x uint64 to the buffer bound at RT_INDEX_DEBUG            */
    uint32_t tileRows;         /* 0: the launch's rows are the contiguous global rows row0..row0+rows-1.
                                  n > 0: interleaved row tiles for multi-GPU load balance: local row r
                                  is global row  row0 + (r / n) * tileStride + (r % n)            */
    uint32_t tileStride;       /* distance in rows between this launch's consecutive tiles (n * ranks) */
} RTLaunchParams;

typedef struct RTDeviceInfo {
    char     name[128];
    int32_t  ordinal;
    int32_t  smCount;
    int32_t  ccMajor, ccMinor;
    int32_t  clockKHz;         /* max SM clock */
    int32_t  supportsRaytracing; /* always 1: the kernel intersects in software
                                    (Sources/gpuRaytracer/main.swift:75 guard) */
    uint64_t totalMemory;
} RTDeviceInfo;

/* ---- errors ---------------------------------------------------------------- */
const char *rtLastErrorString(void);
int         rtAbiVersion(void);

/* ---- device: MTLCreateSystemDefaultDevice()   renderer.swift:30 ------------- */
int rtDeviceCreate(int ordinal, RTDevice **out);
int rtDeviceGetInfo(RTDevice *device, RTDeviceInfo *info);
int rtDeviceRelease(RTDevice *device);

/* ---- pipeline: makeLibrary(URL:) + makeFunction(name:) +
 *      makeComputePipelineState(function:)   renderer.swift:32-43,
 *      Sources/gpuRaytracer/computeShader.swift:112-123.
 *      kernelName is "pathTrace" or "drawTriangle"; anything else is
 *      RT_ERR_UNKNOWN_KERNEL (Metal returns nil and the reference fatalError()s). */
int rtPipelineCreate(RTDevice *device, const char *kernelName, RTPipeline **out);
int rtPipelineRelease(RTPipeline *pipeline);

/* ---- queue: makeCommandQueue()   renderer.swift:44 -------------------------- */
int rtQueueCreate(RTDevice *device, RTQueue **out);
/* Interop: run on an existing CUstream (e.g. torch.cuda.current_stream().cuda_stream). */
int rtQueueCreateWithStream(RTDevice *device, void *cuStream, RTQueue **out);
int rtQueueRelease(RTQueue *queue);

/* ---- buffers: makeBuffer(bytes:length:options:.storageModeShared)
 *      renderer.swift:61-72; .contents()   computeShader.swift:196,203 ---------
 * The bytes are copied (the caller's array may die).  bytes == NULL gives a
 * zero-filled buffer (makeBuffer(length:options:)).  rtBufferContents returns a
 * host pointer valid until release; if a command wrote the buffer since the
 * last call, the device copy is fetched first (after rtWaitUntilCompleted this
 * is what .contents() shows on unified memory).  After writing through the
 * pointer call rtBufferDidModifyRange (MTLBuffer.didModifyRange): the range is copied
 * to the device in stream order behind the last command that used the buffer, without
 * a host-side wait; do not write it again before the next rtBufferContents call
 * (which waits for that stream). */
int      rtBufferCreate(RTDevice *device, const void *bytesOrNull, size_t length, RTBuffer **out);
/* Wrap device memory owned by someone else (a torch tensor); no host mirror. */
int      rtBufferCreateWithDevicePointer(RTDevice *device, uint64_t devicePointer, size_t length, RTBuffer **out);
void    *rtBufferContents(RTBuffer *buffer);
int      rtBufferDidModifyRange(RTBuffer *buffer, size_t offset, size_t length);
size_t   rtBufferLength(RTBuffer *buffer);
uint64_t rtBufferDevicePointer(RTBuffer *buffer);
int      rtBufferRelease(RTBuffer *buffer);

/* ---- textures: makeTexture(descriptor:) renderer.swift:74-93,
 *      replace(region:...) :108-111, getBytes(...) image.swift:35 --------------
 * The kernel uses integer coordinates and no filtering, so a texture is a
 * row-major linear buffer.  RGBA32_FLOAT replaces the reference's rgba16Float
 * render target; R32_UINT is the per-pixel Halton offset. */
int rtTextureCreate(RTDevice *device, int format, int width, int height, RTTexture **out);
int rtTextureReplaceRegion(RTTexture *texture, int x, int y, int width, int height,
                           const void *bytes, size_t bytesPerRow);
int rtTextureGetBytes(RTTexture *texture, void *bytes, size_t bytesPerRow,
                      int x, int y, int width, int height);
RTBuffer *rtTextureBuffer(RTTexture *texture);   /* the storage, still owned by the texture */
int rtTextureRelease(RTTexture *texture);

/* ---- acceleration structure: setupAccelerationStructures()
 *      computeShader.swift:45-97.  Nothing is built: the kernel loops over the
 *      triangles themselves.  The handle records the geometry so that call
 *      sites map one to one and T can default from it. */
int rtAccelerationStructureCreate(RTDevice *device, RTBuffer *vertexBuffer, size_t vertexBufferOffset,
                                  size_t vertexStride, uint32_t triangleCount, RTAccel **out);
int rtAccelerationStructureRelease(RTAccel *accel);

/* ---- encoding: makeCommandBuffer() + makeComputeCommandEncoder() +
 *      setComputePipelineState   renderer.swift:118-121 ------------------------ */
int rtCommandBegin(RTQueue *queue, RTPipeline *pipeline, RTCommand **out);
/* setBuffer(_:offset:index:)   renderer.swift:122-128 */
int rtSetBuffer(RTCommand *command, RTBuffer *buffer, size_t offset, int index);
/* setTexture(_:index:)   renderer.swift:129-130; index 0 -> RT_INDEX_OUTPUT, 1 -> RT_INDEX_OFFSETS */
int rtSetTexture(RTCommand *command, RTTexture *texture, int index);
/* setAccelerationStructure(_:bufferIndex:)   renderer.swift:127 */
int rtSetAccelerationStructure(RTCommand *command, RTAccel *accel, int index);
/* new: the shader's literals as launch parameters */
int rtDefaultParams(int width, int height, RTLaunchParams *out);
int rtSetParams(RTCommand *command, const RTLaunchParams *params);
/* dispatchThreads(_:threadsPerThreadgroup:)   renderer.swift:137-141.  The grid is
 * width x height x 1 (one thread per pixel); the threadgroup shape is advisory,
 * the library launches its own tuned block shape. */
int rtDispatchThreads(RTCommand *command, int gridX, int gridY, int gridZ,
                      int threadsX, int threadsY, int threadsZ);
/* endEncoding() + commit()   renderer.swift:142-143 (asynchronous) */
int rtCommit(RTCommand *command);
/* waitUntilCompleted()   renderer.swift:144 */
int rtWaitUntilCompleted(RTCommand *command);
/* MTLCommandBuffer.gpuStartTime / gpuEndTime: device time of the dispatch, CUDA events */
int rtCommandGetGpuTimeMs(RTCommand *command, double *milliseconds);
/* number of kernels this command launched (prep + trace) */
int rtCommandGetLaunchCount(RTCommand *command, int *count);
int rtCommandRelease(RTCommand *command);

/* ---- resolve / image output: saveTextureToImage   RTrace/image.swift:40-65 ----
 * in: float4 per pixel; out: RGBA8.  divideByAlpha != 0 resolves a progressive
 * accumulation buffer (xyz / w) first.  Runs as a small kernel on the queue. */
int rtResolveToRGBA8(RTQueue *queue, RTBuffer *in, size_t inOffset, RTBuffer *out, size_t outOffset,
                     uint64_t pixelCount, int divideByAlpha);
/* The same buffer in the reference's render-target format, rgba16Float (renderer.swift:74-82): 8 bytes per pixel,
 * each channel rounded to nearest-even as the shader's texture write does -- what saveTextureToImage reads back as
 * [Float16] (image.swift:25-38).  divideByAlpha != 0: xyz / w first, alpha 1. */
int rtResolveToRGBA16F(RTQueue *queue, RTBuffer *in, size_t inOffset, RTBuffer *out, size_t outOffset,
                       uint64_t pixelCount, int divideByAlpha);

/* ---- pathTrace without any allocation ----------------------------------------
 * No counterpart in the reference; added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Renders as a pathTrace command does (same parameters, same image bit for bit) but calls no cuMemAlloc / cudaMalloc /
 * cudaFree nor any allocating library routine: the kernel's scratch -- the prepared scene, rt_scene_layout.h -- is a
 * static __device__ array of 57 664 bytes inside the library (every GPU that loads it holds it until exit), and the
 * Halton values are computed instead of read from the library's tables.
 *   - all pointers are DEVICE pointers (rtBufferDevicePointer of a library buffer, a torch tensor's data_ptr(), ...):
 *     camera CameraGPU[1], materials MaterialGPU[T], lights SquareLightGPU[L], vertices simd_float3[3T], spheres
 *     SphereGPU[S] or NULL, offsets uint32[width * rows], out float4[width * rows] (this launch's rows);
 *   - limits: numTriangles <= 256, numSpheres <= 256, numLights <= 16, else RT_ERR_OUT_OF_RANGE and nothing is
 *     launched; params->variant may carry RT_VARIANT_LIGHT_EXTENT only (there is one kernel form);
 *   - the device to render on must be CURRENT on the calling thread (cudaSetDevice / an rtDeviceCreate'd device's
 *     primary context); cuStream is a CUstream / cudaStream_t of that device or NULL; the call is asynchronous;
 *   - every call on a device shares the one scratch array, which keeps the last call's contents (each call clears
 *     the part it uses first): calls that OVERLAP on the GPU -- two streams, or a captured graph replayed beside a
 *     live call -- race on it.  Serialise them (one stream, or events).
 * Returns RT_OK, an RT_ERR_* code, or > 0: a cudaError_t of the CUDA runtime. */
int rtPathTraceNoAlloc(const void *camera, const void *materials, const void *lights, const void *vertices,
                       const void *spheresOrNull, const uint32_t *offsets, float *out,
                       const RTLaunchParams *params, void *cuStream);

/* ---- new: one process, several GPUs, row tiles (SURVEY 8b "Multi-GPU (new)", 8e) ----
 * rtGroupCreate makes one device + queue per ordinal (an ordinal may repeat) and lets the members reach member 0's
 * memory peer to peer where the hardware allows it. */
int rtGroupCreate(const int *ordinals, int count, RTGroup **out);
int rtGroupDeviceCount(RTGroup *group);
/* Renders the frame of params->width x params->height with pathTrace: the rows are dealt to the members in INTERLEAVED
 * tiles of 8 rows (tile k -> member k mod n; the kernel maps a member's compact rows to global rows, so indexing is a
 * one-GPU render's and the image is the same bit for bit), all members get the same scene bytes, each copies its tiles
 * device to device straight into place in member 0's frame buffer on its own stream, and member 0 reads the frame back
 * ONCE into hostOut (width*height*16 bytes).  params->row0 / rows / tileRows / tileStride are set by the library.
 * params->accumulate != 0 continues the running sums in hostOut.  Pipelines and buffers are made on the first call and
 * reused.  gpuMs (may be NULL) receives the slowest member's device time.  Not thread-safe per group. */
int rtGroupRender(RTGroup *group, const char *kernelName,
                  const void *camera, const void *materials, size_t materialsLength,
                  const void *lights, size_t lightsLength,
                  const void *vertices, size_t verticesLength,
                  const void *spheres, size_t spheresLength,
                  const uint32_t *offsets, const RTLaunchParams *params,
                  float *hostOut, double *gpuMs);
int rtGroupRelease(RTGroup *group);

#ifdef __cplusplus
}
#endif
#endif /* RTCUDA_H */
