/*
 * noalloc.cu -- rtPathTraceNoAlloc (include/rtcuda.h): pathTrace for callers that may not allocate device memory
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 *
 * No counterpart in the reference.
 *
 * Every other path through the library keeps its scratch -- the PREPARED SCENE that rt_prep_scene writes and the
 * tracer stages into shared memory (rt_scene_layout.h), the Halton tables, the ticket flags -- in memory it
 * cuMemAlloc()s per queue or device.  This entry point allocates nothing: its prepared scene lives in the file-scope
 * __device__ array below, RT_NOALLOC_SCENE_WORDS words (57.7 KB: up to 256 triangles, 256 spheres and 16 lights;
 * every GPU that loads the library's runtime module holds it until the process exits), and its Halton values are
 * computed instead of looked up (the tracer is the compacting kernel of wave.cuh with an empty table: the same
 * arithmetic, the same image bit for bit).  The caller passes every input and the output as device pointers.
 *
 * In HOST code the array's name is a host-side stand-in, not a device address, so its device address is fetched
 * once per device with cudaGetSymbolAddress (the device current, the return code checked) and handed to the kernels
 * as an ordinary pointer.
 *
 * This is the one translation unit of the library on the CUDA RUNTIME API (cudaGetSymbolAddress is a runtime call;
 * libcudart is linked statically and loads the driver lazily, so the library still loads without one).  It shares the
 * device's primary context with rtcuda.cpp's driver-API objects, so buffers made by rtBufferCreate can be passed here
 * by rtBufferDevicePointer.
 *
 * The kernels are pathtrace.cu's own (rt_prep_scene, pathtrace_wave_body), compiled a second time into this module
 * with the same flags (--fmad=false -prec-div=true -prec-sqrt=true -ftz=false): RT_NOALLOC_BUILD leaves out
 * every other entry of that file.
 */
#define RT_NOALLOC_BUILD
#define rt_prep_scene rt_noalloc_prep_scene      /* this module's copy of the prep kernel: keep the symbol apart */
#include "pathtrace.cu"
#undef rt_prep_scene

#include <stdarg.h>
#include <stdio.h>

/* the largest scene this entry point takes, and the prepared scene's size for it (rt_scene_layout.h) */
#define RT_NOALLOC_MAX_TRIANGLES 256u
#define RT_NOALLOC_MAX_SPHERES   256u
#define RT_NOALLOC_MAX_LIGHTS    16u
#define RT_NOALLOC_SCENE_WORDS   (RT_HDR_WORDS + (RT_ISECT_WORDS + RT_SHADE_ROWS) * RT_NOALLOC_MAX_TRIANGLES + \
                                  (RT_SPHERE_WORDS + RT_SSHADE_ROWS) * RT_NOALLOC_MAX_SPHERES + RT_LIGHT_WORDS * RT_NOALLOC_MAX_LIGHTS + \
                                  RT_GROUP_WORDS + 4)

/* THE scratch: one per device, shared by every call; it keeps the last call's contents (each call clears the words it
 * uses first), and two calls that overlap on the GPU -- different streams, or a graph replayed beside a live call --
 * race on it. */
static __device__ __align__(256) float g_noalloc_scene[RT_NOALLOC_SCENE_WORDS];

static __global__ void __launch_bounds__(RT_BLOCK_THREADS, 4)
pathTrace_noalloc6(const float *__restrict__ scene, const uint32_t *__restrict__ offsets, float4 *__restrict__ out,
                   const RTLaunchParams p, const RTSceneLayout lay) {
    extern __shared__ __align__(16) float sm[];
    pathtrace_wave_body<false, 4, 6, false, false>(scene, offsets, out, p, lay, sm, nullptr, RTTileSched{}, RTHaltonTables{});
}
/* R = 4: for scenes whose prepared form leaves no room for four R = 6 blocks per SM (rt_plan.cpp has the same rule) */
static __global__ void __launch_bounds__(RT_BLOCK_THREADS, 4)
pathTrace_noalloc4(const float *__restrict__ scene, const uint32_t *__restrict__ offsets, float4 *__restrict__ out,
                   const RTLaunchParams p, const RTSceneLayout lay) {
    extern __shared__ __align__(16) float sm[];
    pathtrace_wave_body<false, 4, 4, false, false>(scene, offsets, out, p, lay, sm, nullptr, RTTileSched{}, RTHaltonTables{});
}

extern "C" void rtInternalSetError(const char *text);       /* rtcuda.cpp: what rtLastErrorString() returns on this thread */

static int naFail(int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    rtInternalSetError(buf);
    return code;
}
static int naCuda(cudaError_t e, const char *what) {
    return naFail((int)e, "rtPathTraceNoAlloc: %s failed: CUDA runtime error %d (%s)", what, (int)e, cudaGetErrorString(e));
}

/* set-up, once per device: the scratch array's DEVICE address, the kernels' shared-memory limit, the SM's capacity */
#define RT_NOALLOC_MAX_DEVICES 64
struct NoAllocDevice { int ready; float *scene; size_t maxSmemOptin, smemPerSM; };
static NoAllocDevice g_na[RT_NOALLOC_MAX_DEVICES];

static int naSetup(int dev, NoAllocDevice **out) {
    if (dev < 0 || dev >= RT_NOALLOC_MAX_DEVICES) return naFail(RT_ERR_UNSUPPORTED, "rtPathTraceNoAlloc: device ordinal %d beyond %d", dev, RT_NOALLOC_MAX_DEVICES);
    NoAllocDevice *d = &g_na[dev];
    if (!__atomic_load_n(&d->ready, __ATOMIC_ACQUIRE)) {
        float *ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void **)&ptr, g_noalloc_scene);      /* the current device's copy of the array */
        if (e != cudaSuccess) return naCuda(e, "cudaGetSymbolAddress(g_noalloc_scene)");
        if (!ptr) return naFail(RT_ERR_STATE, "rtPathTraceNoAlloc: cudaGetSymbolAddress returned a null address");
        int optin = 0, perSM = 0;
        if ((e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev)) != cudaSuccess) return naCuda(e, "cudaDeviceGetAttribute");
        if ((e = cudaDeviceGetAttribute(&perSM, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev)) != cudaSuccess) return naCuda(e, "cudaDeviceGetAttribute");
        if ((e = cudaFuncSetAttribute(pathTrace_noalloc6, cudaFuncAttributeMaxDynamicSharedMemorySize, optin - 1024)) != cudaSuccess) return naCuda(e, "cudaFuncSetAttribute");
        if ((e = cudaFuncSetAttribute(pathTrace_noalloc4, cudaFuncAttributeMaxDynamicSharedMemorySize, optin - 1024)) != cudaSuccess) return naCuda(e, "cudaFuncSetAttribute");
        d->scene = ptr; d->maxSmemOptin = (size_t)optin; d->smemPerSM = (size_t)perSM - 1024;
        __atomic_store_n(&d->ready, 1, __ATOMIC_RELEASE);
    }
    *out = d;
    return RT_OK;
}

extern "C" int rtPathTraceNoAlloc(const void *camera, const void *materials, const void *lights, const void *vertices,
                                  const void *spheresOrNull, const uint32_t *offsets, float *out,
                                  const RTLaunchParams *params, void *cuStream) {
    if (!camera || !materials || !lights || !vertices || !offsets || !out || !params)
        return naFail(RT_ERR_INVALID_ARGUMENT, "rtPathTraceNoAlloc: NULL argument");
    RTLaunchParams p = *params;
    if (p.numLights == 0) p.numLights = 1;
    if (p.width <= 0 || p.height <= 0 || p.row0 < 0 || p.rows <= 0 || p.spp == 0)
        return naFail(RT_ERR_INVALID_ARGUMENT, "rtPathTraceNoAlloc: bad image / tile size or spp = 0");
    {
        const long last = p.tileRows ? (long)p.row0 + (long)((p.rows - 1) / p.tileRows) * p.tileStride + (p.rows - 1) % p.tileRows
                                     : (long)p.row0 + p.rows - 1;
        if (last >= p.height || (p.tileRows && p.tileStride < p.tileRows))
            return naFail(RT_ERR_INVALID_ARGUMENT, "rtPathTraceNoAlloc: last row %ld of %d", last, p.height);
    }
    if (p.bounces < 1 || p.bounces > 4)
        return naFail(RT_ERR_UNSUPPORTED, "bounces = %d; Halton dimension 2+5b+3 must stay inside primes[24] (1..4)", p.bounces);
    if (p.variant & ~(uint32_t)RT_VARIANT_LIGHT_EXTENT)      /* (RT_VARIANT_MIS included: this entry point has the reference's estimator only) */
        return naFail(RT_ERR_UNSUPPORTED, "rtPathTraceNoAlloc has one kernel form: variant may carry RT_VARIANT_LIGHT_EXTENT only");
    if (p.numTriangles == 0 || p.numTriangles > RT_NOALLOC_MAX_TRIANGLES || p.numSpheres > RT_NOALLOC_MAX_SPHERES || p.numLights > RT_NOALLOC_MAX_LIGHTS)
        return naFail(RT_ERR_OUT_OF_RANGE, "rtPathTraceNoAlloc: scene of %u triangles, %u spheres, %u lights exceeds the static scratch "
                      "(at most %u triangles, %u spheres, %u lights; use rtCommit for larger scenes)", p.numTriangles, p.numSpheres, p.numLights,
                      RT_NOALLOC_MAX_TRIANGLES, RT_NOALLOC_MAX_SPHERES, RT_NOALLOC_MAX_LIGHTS);
    if (p.numSpheres && !spheresOrNull) return naFail(RT_ERR_MISSING_BINDING, "rtPathTraceNoAlloc: numSpheres = %u without a sphere buffer", p.numSpheres);

    RTSceneLayout lay = rtSceneLayout(p.numTriangles, p.numSpheres, p.numLights);
    if (p.variant & RT_VARIANT_LIGHT_EXTENT) lay.flags |= RT_LAYOUT_LIGHT_EXTENT;
    lay.flags |= RT_LAYOUT_NO_SGRID;            /* the static scratch holds the staged scene only: spheres by brute force */
    if (lay.words > RT_NOALLOC_SCENE_WORDS)     /* (cannot happen within the three limits above; the array's size is what decides) */
        return naFail(RT_ERR_OUT_OF_RANGE, "rtPathTraceNoAlloc: prepared scene of %u words exceeds the static scratch of %u", lay.words, (unsigned)RT_NOALLOC_SCENE_WORDS);

    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return naCuda(e, "cudaGetDevice");
    NoAllocDevice *d = nullptr;
    int rc = naSetup(dev, &d);
    if (rc) return rc;

    const size_t sceneBytes = (size_t)lay.words * 4;
    const size_t wave6 = (size_t)rtWaveWords(6, lay.T, lay.S) * 4, wave4 = (size_t)rtWaveWords(4, lay.T, lay.S) * 4;
    const bool r6 = d->smemPerSM / (sceneBytes + wave6 + 1024) >= 4;
    const size_t smem = sceneBytes + (r6 ? wave6 : wave4);
    if (smem + 1024 > d->maxSmemOptin) return naFail(RT_ERR_UNSUPPORTED, "rtPathTraceNoAlloc: scene + path queues (%zu bytes) do not fit shared memory", smem);

    cudaStream_t stream = (cudaStream_t)cuStream;
    /* the array keeps the last call's contents: clear what this call uses, then prep rewrites every word the tracer reads */
    if ((e = cudaMemsetAsync(d->scene, 0, sceneBytes, stream)) != cudaSuccess) return naCuda(e, "cudaMemsetAsync");
    rt_noalloc_prep_scene<<<1, 256, 0, stream>>>((const unsigned char *)camera, (const unsigned char *)materials, (const unsigned char *)lights,
                                                 (const unsigned char *)vertices, (const unsigned char *)spheresOrNull, d->scene, lay,
                                                 (const float4 *)nullptr);
    if ((e = cudaGetLastError()) != cudaSuccess) return naCuda(e, "launch of rt_prep_scene");
    const dim3 grid((unsigned)((p.width + 15) / 16), (unsigned)((p.rows + 7) / 8));
    if (r6) pathTrace_noalloc6<<<grid, RT_BLOCK_THREADS, smem, stream>>>(d->scene, offsets, (float4 *)out, p, lay);
    else    pathTrace_noalloc4<<<grid, RT_BLOCK_THREADS, smem, stream>>>(d->scene, offsets, (float4 *)out, p, lay);
    if ((e = cudaGetLastError()) != cudaSuccess) return naCuda(e, "launch of pathTrace");
    return RT_OK;
}
