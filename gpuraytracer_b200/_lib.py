"""ctypes binding of librtcuda.so (include/rtcuda.h).

The shared library is built in-tree by ``__graft_entry__.build()`` /
``make -C gpuraytracer_b200/csrc``.  There is no Python or CPU fallback: if the
library is missing, importing this module raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RTCUDA_LIB") or os.path.join(_HERE, "librtcuda.so")   # (override: kernel A/B experiments)

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(there is no fallback implementation)")

lib = C.CDLL(LIB_PATH)

RT_OK = 0
RT_ERR = {-1: "INVALID_ARGUMENT", -2: "UNKNOWN_KERNEL", -3: "NO_DRIVER", -4: "MISSING_BINDING",
          -5: "OUT_OF_RANGE", -6: "STATE", -7: "UNSUPPORTED", -8: "OUT_OF_MEMORY", -9: "DEVICE_TIMEOUT"}

INDEX_CAMERA, INDEX_MATERIALS, INDEX_LIGHTS, INDEX_VERTICES, INDEX_PIXELS = 0, 1, 2, 3, 4
INDEX_ACCEL, INDEX_DEBUG, INDEX_SPHERES, INDEX_OUTPUT, INDEX_OFFSETS = 5, 6, 7, 16, 17
INDEX_LIGHT_NORMALS = 8        # optional float4[L] side table: the lights' normals
FORMAT_RGBA32_FLOAT, FORMAT_R32_UINT = 1, 2
VARIANT_DEFAULT, VARIANT_SMEM, VARIANT_CONST, VARIANT_WAVE, VARIANT_TILES, VARIANT_WIDE = 0, 1, 2, 3, 16, 17
VARIANT_SORTED, VARIANT_SORTED_TILES, VARIANT_SORTED_BLOCK, VARIANT_SORTED_BLOCK_TILES = 18, 19, 20, 21
VARIANT_BVH = 22
VARIANT_COUNT_RAYS = 0x100
VARIANT_LIGHT_EXTENT = 0x200   # flag: sampleAreaLight over the light's own width x depth
VARIANT_MIS = 0x400            # flag: light sample + cosine-sampled continuation combined by the balance heuristic


class RTLaunchParams(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("row0", C.c_int32), ("rows", C.c_int32),
                ("spp", C.c_uint32), ("firstSample", C.c_uint32), ("bounces", C.c_int32),
                ("accumulate", C.c_int32), ("numTriangles", C.c_uint32), ("numSpheres", C.c_uint32),
                ("numLights", C.c_uint32), ("cameraRays", C.c_uint32), ("misSamples", C.c_uint32),
                ("variant", C.c_uint32), ("tileRows", C.c_uint32), ("tileStride", C.c_uint32)]


class RTDeviceInfo(C.Structure):
    _fields_ = [("name", C.c_char * 128), ("ordinal", C.c_int32), ("smCount", C.c_int32),
                ("ccMajor", C.c_int32), ("ccMinor", C.c_int32), ("clockKHz", C.c_int32),
                ("supportsRaytracing", C.c_int32), ("totalMemory", C.c_uint64)]


class RTHostSceneBytes(C.Structure):
    _fields_ = [("camera", C.c_void_p), ("materials", C.c_void_p), ("capTriangles", C.c_uint32),
                ("vertices", C.c_void_p), ("lights", C.c_void_p), ("capLights", C.c_uint32),
                ("spheres", C.c_void_p), ("capSpheres", C.c_uint32),
                ("numTriangles", C.c_uint32), ("numSpheres", C.c_uint32), ("numLights", C.c_uint32)]


P = C.c_void_p
PP = C.POINTER(C.c_void_p)
# name -> (restype, argtypes); every symbol include/rtcuda.h declares
SIGNATURES = {
    "rtLastErrorString": (C.c_char_p, []),
    "rtAbiVersion": (C.c_int, []),
    "rtDeviceCreate": (C.c_int, [C.c_int, PP]),
    "rtDeviceGetInfo": (C.c_int, [P, C.POINTER(RTDeviceInfo)]),
    "rtDeviceRelease": (C.c_int, [P]),
    "rtPipelineCreate": (C.c_int, [P, C.c_char_p, PP]),
    "rtPipelineRelease": (C.c_int, [P]),
    "rtQueueCreate": (C.c_int, [P, PP]),
    "rtQueueCreateWithStream": (C.c_int, [P, P, PP]),
    "rtQueueRelease": (C.c_int, [P]),
    "rtBufferCreate": (C.c_int, [P, P, C.c_size_t, PP]),
    "rtBufferCreateWithDevicePointer": (C.c_int, [P, C.c_uint64, C.c_size_t, PP]),
    "rtBufferContents": (C.c_void_p, [P]),
    "rtBufferDidModifyRange": (C.c_int, [P, C.c_size_t, C.c_size_t]),
    "rtBufferLength": (C.c_size_t, [P]),
    "rtBufferDevicePointer": (C.c_uint64, [P]),
    "rtBufferRelease": (C.c_int, [P]),
    "rtTextureCreate": (C.c_int, [P, C.c_int, C.c_int, C.c_int, PP]),
    "rtTextureReplaceRegion": (C.c_int, [P, C.c_int, C.c_int, C.c_int, C.c_int, P, C.c_size_t]),
    "rtTextureGetBytes": (C.c_int, [P, P, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int]),
    "rtTextureBuffer": (C.c_void_p, [P]),
    "rtTextureRelease": (C.c_int, [P]),
    "rtAccelerationStructureCreate": (C.c_int, [P, P, C.c_size_t, C.c_size_t, C.c_uint32, PP]),
    "rtAccelerationStructureRelease": (C.c_int, [P]),
    "rtCommandBegin": (C.c_int, [P, P, PP]),
    "rtSetBuffer": (C.c_int, [P, P, C.c_size_t, C.c_int]),
    "rtSetTexture": (C.c_int, [P, P, C.c_int]),
    "rtSetAccelerationStructure": (C.c_int, [P, P, C.c_int]),
    "rtDefaultParams": (C.c_int, [C.c_int, C.c_int, C.POINTER(RTLaunchParams)]),
    "rtSetParams": (C.c_int, [P, C.POINTER(RTLaunchParams)]),
    "rtDispatchThreads": (C.c_int, [P] + [C.c_int] * 6),
    "rtCommit": (C.c_int, [P]),
    "rtWaitUntilCompleted": (C.c_int, [P]),
    "rtCommandGetGpuTimeMs": (C.c_int, [P, C.POINTER(C.c_double)]),
    "rtCommandGetLaunchCount": (C.c_int, [P, C.POINTER(C.c_int)]),
    "rtCommandRelease": (C.c_int, [P]),
    "rtResolveToRGBA8": (C.c_int, [P, P, C.c_size_t, P, C.c_size_t, C.c_uint64, C.c_int]),
    "rtResolveToRGBA16F": (C.c_int, [P, P, C.c_size_t, P, C.c_size_t, C.c_uint64, C.c_int]),
    "rtPathTraceNoAlloc": (C.c_int, [P, P, P, P, P, P, P, C.POINTER(RTLaunchParams), P]),
    "rtGroupCreate": (C.c_int, [C.POINTER(C.c_int), C.c_int, PP]),
    "rtGroupDeviceCount": (C.c_int, [P]),
    "rtGroupRender": (C.c_int, [P, C.c_char_p, P, P, C.c_size_t, P, C.c_size_t, P, C.c_size_t, P, C.c_size_t,
                                P, C.POINTER(RTLaunchParams), P, C.POINTER(C.c_double)]),
    "rtGroupRelease": (C.c_int, [P]),
}
# host-side mirror + test hook (not part of rtcuda.h)
EXTRA_SIGNATURES = {
    "rtHostInitCornellBox": (C.c_int, [C.c_int, C.c_int, C.c_float, C.POINTER(RTHostSceneBytes)]),
    "rtHostInitSphereScene": (C.c_int, [C.c_int, C.c_int, C.c_uint32, C.c_uint32, C.POINTER(RTHostSceneBytes)]),
    "rtHostFillOffsets": (None, [P, C.c_int, C.c_int, C.c_int, C.c_uint32]),
    "rtDebugEvaluate": (C.c_int, [P, C.c_int, P, C.c_size_t, P, C.c_size_t, C.c_uint32]),
    "rtDebugSetTuning": (C.c_int, [P, C.c_char_p, C.c_int]),
    "rtDebugMeasureFp32Peak": (C.c_int, [P, C.POINTER(C.c_double)]),
    "rtDebugChooseKernel": (C.c_int, [C.POINTER(RTLaunchParams), C.c_int, C.c_int, C.c_size_t, C.c_size_t, C.c_char_p, C.c_size_t,
                                      C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
}
for _name, (_res, _args) in {**SIGNATURES, **EXTRA_SIGNATURES}.items():
    _f = getattr(lib, _name)          # AttributeError here = the library lacks a declared symbol
    _f.restype = _res
    _f.argtypes = _args


class RTError(RuntimeError):
    def __init__(self, code, where):
        self.code = code
        text = lib.rtLastErrorString().decode(errors="replace")
        kind = RT_ERR.get(code, f"CUresult {code}" if code > 0 else str(code))
        super().__init__(f"{where}: {kind}: {text}")


def check(code, where):
    if code != RT_OK:
        raise RTError(code, where)
