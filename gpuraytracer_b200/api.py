"""Python objects over the C ABI, one per Metal object the reference host uses
(RTrace/renderer.swift:9-27), plus ``Renderer``: the mirror of the reference's
``Renderer.init`` / ``Renderer.draw`` (renderer.swift:29-146).

Everything here ends in a call into librtcuda.so; nothing is computed in Python.
"""
import ctypes as C

import numpy as np

from . import _lib as L
from ._lib import RTLaunchParams, check, lib


def _vp(a):
    return a.ctypes.data_as(C.c_void_p)


class _Handle:
    _release = None

    def __init__(self, handle):
        self.h = handle

    def release(self):
        if getattr(self, "h", None):
            self._release(self.h)
            self.h = None

    def __del__(self):
        try:
            self.release()
        except Exception:
            pass


class Device(_Handle):
    """MTLCreateSystemDefaultDevice(), renderer.swift:30."""
    _release = staticmethod(lib.rtDeviceRelease)

    def __init__(self, ordinal=0):
        h = C.c_void_p()
        check(lib.rtDeviceCreate(ordinal, C.byref(h)), "rtDeviceCreate")
        super().__init__(h)

    def info(self):
        i = L.RTDeviceInfo()
        check(lib.rtDeviceGetInfo(self.h, C.byref(i)), "rtDeviceGetInfo")
        return i

    def set_tuning(self, key, value):
        """A/B knob of this device (csrc/rt_plan.h RTTuning; not part of rtcuda.h): tests and profiling only."""
        check(lib.rtDebugSetTuning(self.h, key.encode(), int(value)), "rtDebugSetTuning")

    def measure_fp32_peak(self):
        """(FFMA, FFMA2) TFLOP/s of an FMA-only kernel on this device (csrc/pathtrace.cu: rt_fp32_peak), ~10 ms."""
        t = (C.c_double * 2)()
        check(lib.rtDebugMeasureFp32Peak(self.h, t), "rtDebugMeasureFp32Peak")
        return t[0], t[1]

    def make_buffer(self, data=None, length=None):
        return Buffer(self, data, length)

    def make_pipeline(self, name):
        return Pipeline(self, name)

    def make_queue(self, stream=None):
        return Queue(self, stream)


class Pipeline(_Handle):
    """makeFunction(name:) + makeComputePipelineState, renderer.swift:39-43."""
    _release = staticmethod(lib.rtPipelineRelease)

    def __init__(self, device, name):
        h = C.c_void_p()
        check(lib.rtPipelineCreate(device.h, name.encode(), C.byref(h)), "rtPipelineCreate")
        super().__init__(h)
        self.device = device


class Queue(_Handle):
    """makeCommandQueue(), renderer.swift:44."""
    _release = staticmethod(lib.rtQueueRelease)

    def __init__(self, device, stream=None):
        h = C.c_void_p()
        if stream is None:
            check(lib.rtQueueCreate(device.h, C.byref(h)), "rtQueueCreate")
        else:
            check(lib.rtQueueCreateWithStream(device.h, C.c_void_p(stream), C.byref(h)), "rtQueueCreateWithStream")
        super().__init__(h)
        self.device = device


class Buffer(_Handle):
    """makeBuffer(bytes:length:options:), renderer.swift:61-72."""
    _release = staticmethod(lib.rtBufferRelease)

    def __init__(self, device, data=None, length=None, device_pointer=None):
        h = C.c_void_p()
        if device_pointer is not None:
            check(lib.rtBufferCreateWithDevicePointer(device.h, device_pointer, length, C.byref(h)),
                  "rtBufferCreateWithDevicePointer")
        elif data is not None:
            a = np.ascontiguousarray(data)
            check(lib.rtBufferCreate(device.h, _vp(a), a.nbytes, C.byref(h)), "rtBufferCreate")
            length = a.nbytes
        else:
            check(lib.rtBufferCreate(device.h, None, length, C.byref(h)), "rtBufferCreate")
        super().__init__(h)
        self.device, self.length = device, length

    def contents(self, dtype=np.uint8):
        """MTLBuffer.contents(): a numpy view of the host-visible bytes."""
        p = lib.rtBufferContents(self.h)
        if not p:
            raise L.RTError(-6, "rtBufferContents")
        raw = (C.c_uint8 * self.length).from_address(p)
        return np.frombuffer(raw, dtype=dtype)

    def did_modify(self, offset=0, length=None):
        check(lib.rtBufferDidModifyRange(self.h, offset, self.length - offset if length is None else length),
              "rtBufferDidModifyRange")

    @property
    def device_pointer(self):
        return lib.rtBufferDevicePointer(self.h)


class Command(_Handle):
    """makeCommandBuffer + makeComputeCommandEncoder, renderer.swift:118-144."""
    _release = staticmethod(lib.rtCommandRelease)

    def __init__(self, queue, pipeline):
        h = C.c_void_p()
        check(lib.rtCommandBegin(queue.h, pipeline.h, C.byref(h)), "rtCommandBegin")
        super().__init__(h)
        self._keep = [queue, pipeline]

    def set_buffer(self, buffer, offset, index):
        check(lib.rtSetBuffer(self.h, buffer.h, offset, index), "rtSetBuffer")
        self._keep.append(buffer)

    def set_params(self, params):
        check(lib.rtSetParams(self.h, C.byref(params)), "rtSetParams")

    def dispatch_threads(self, grid, threads=(8, 8, 1)):
        check(lib.rtDispatchThreads(self.h, grid[0], grid[1], grid[2], *threads), "rtDispatchThreads")

    def commit(self):
        check(lib.rtCommit(self.h), "rtCommit")

    def wait_until_completed(self):
        check(lib.rtWaitUntilCompleted(self.h), "rtWaitUntilCompleted")

    def gpu_time_ms(self):
        ms = C.c_double()
        check(lib.rtCommandGetGpuTimeMs(self.h, C.byref(ms)), "rtCommandGetGpuTimeMs")
        return ms.value

    def launch_count(self):
        n = C.c_int()
        check(lib.rtCommandGetLaunchCount(self.h, C.byref(n)), "rtCommandGetLaunchCount")
        return n.value


B200_SHAPE = dict(sm_count=148, max_smem_optin=232448, smem_per_sm=233472)


def choose_kernel(params, kind=0, sm_count=148, max_smem_optin=232448, smem_per_sm=233472):
    """The kernel a command with these parameters would run on a device of this shape (csrc/rt_plan.cpp), without a GPU.
    Returns dict(kernel, smem_bytes, grid, chunks, chunk_spp)."""
    name = C.create_string_buffer(64)
    smem, grid, ch = C.c_uint32(), (C.c_uint32 * 2)(), (C.c_uint32 * 2)()
    check(lib.rtDebugChooseKernel(C.byref(params), kind, sm_count, max_smem_optin, smem_per_sm, name, 64, C.byref(smem), grid, ch),
          "rtDebugChooseKernel")
    return dict(kernel=name.value.decode(), smem_bytes=smem.value, grid=(grid[0], grid[1]), chunks=ch[0], chunk_spp=ch[1])


def default_params(width, height, **over):
    p = RTLaunchParams()
    check(lib.rtDefaultParams(width, height, C.byref(p)), "rtDefaultParams")
    for k, v in over.items():
        setattr(p, k, v)
    return p
