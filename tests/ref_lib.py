"""ctypes access to oracle/_ref/libref_metal.so: the reference's UNMODIFIED .metal shaders
compiled as C++ (oracle/ref_driver_*.cpp + oracle/metal_compat).  Test infrastructure.
The library can only be BUILT where /root/reference exists; the prebuilt file travels."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATH = os.path.join(ROOT, "oracle", "_ref", "libref_metal.so")


def load():
    if not os.path.exists(PATH) and os.path.isdir("/root/reference/RTrace"):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "ref"])
    if not os.path.exists(PATH):
        return None
    lib = C.CDLL(PATH)
    lib.rt_ref_halton.restype = C.c_float
    lib.rt_ref_halton.argtypes = [C.c_uint32, C.c_uint32]
    lib.rt_ref_hash.restype = C.c_uint32
    lib.rt_ref_hash.argtypes = [C.c_uint32]
    return lib


def vp(a):
    return a.ctypes.data_as(C.c_void_p)


def pathtrace(lib, scene, width, height, offsets, row0=0, rows=None):
    """The reference kernel at its own literals (400 spp, 3 bounces).  offsets: uint32 [height*width]."""
    rows = height - row0 if rows is None else rows
    out = np.zeros((height, width, 4), np.float32)
    rc = lib.rt_ref_pathtrace(vp(scene["camera"]), vp(scene["materials"]), vp(scene["lights"]), vp(scene["vertices"]),
                              C.c_uint32(scene["T"]), vp(offsets), vp(out), width, height, row0, rows)
    assert rc == 0
    return out[row0:row0 + rows]


def drawtriangle(lib, scene, width, height, row0=0, rows=None):
    """The reference MIS kernel at its own literals (6 camera rays, 300 MIS samples)."""
    rows = height - row0 if rows is None else rows
    pixels = np.zeros((height, width, 4), np.uint8)
    text = np.zeros((height, width, 3), np.float32)
    rc = lib.rt_ref_drawtriangle(vp(scene["camera"]), vp(scene["materials"]), vp(scene["lights"]), vp(scene["vertices"]),
                                 C.c_uint32(scene["T"]), vp(pixels), vp(text), width, height, row0, rows)
    assert rc == 0
    return pixels[row0:row0 + rows], text[row0:row0 + rows]
