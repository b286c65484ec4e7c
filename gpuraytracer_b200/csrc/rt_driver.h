/*
 * rt_driver.h -- the CUDA driver API, loaded at run time.
 *
 * librtcuda.so does not link libcuda: it dlopen()s libcuda.so.1 on the first
 * rtDeviceCreate so that the library itself loads (and its symbol table can be
 * checked) on a machine without a GPU driver, and fails loudly with
 * RT_ERR_NO_DRIVER instead of falling back to anything.
 */
#pragma once
#include <cuda.h>

#define RT_STR2(x) #x
#define RT_STR(x) RT_STR2(x)

/* X(name): `name` is the cuda.h macro/identifier, so versioned entry points
 * (cuMemAlloc -> cuMemAlloc_v2 ...) resolve to what the header was written for. */
#define RT_DRIVER_FUNCTIONS(X) \
    X(cuInit) X(cuGetErrorString) X(cuDeviceGet) X(cuDeviceGetCount) X(cuDeviceGetName) \
    X(cuDeviceGetAttribute) X(cuDeviceTotalMem) X(cuDevicePrimaryCtxRetain) \
    X(cuDevicePrimaryCtxRelease) X(cuCtxSetCurrent) X(cuCtxSynchronize) X(cuModuleLoadData) X(cuModuleUnload) \
    X(cuModuleGetFunction) X(cuModuleGetGlobal) X(cuFuncSetAttribute) X(cuMemAlloc) X(cuMemFree) \
    X(cuMemHostAlloc) X(cuMemHostGetDevicePointer) X(cuMemFreeHost) X(cuMemcpyHtoD) X(cuMemcpyDtoH) X(cuMemcpyHtoDAsync) \
    X(cuMemcpyDtoHAsync) X(cuMemcpyDtoDAsync) X(cuMemcpy2DAsync) X(cuCtxEnablePeerAccess) X(cuDeviceCanAccessPeer) X(cuMemsetD8) X(cuMemsetD8Async) X(cuStreamCreate) \
    X(cuStreamDestroy) X(cuStreamSynchronize) X(cuEventCreate) X(cuEventRecord) X(cuStreamWaitEvent) \
    X(cuEventSynchronize) X(cuEventElapsedTime) X(cuEventDestroy) X(cuLaunchKernel) \
    X(cuOccupancyMaxActiveBlocksPerMultiprocessor)

struct RTDriver {
#define RT_DECL(name) decltype(&name) p_##name;
    RT_DRIVER_FUNCTIONS(RT_DECL)
#undef RT_DECL
    void *handle;
};

/* returns NULL (and sets the thread's error text) when libcuda cannot be loaded */
const RTDriver *rtDriver(void);
