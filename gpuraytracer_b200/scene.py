"""Scene bytes from the C++ host mirror (csrc/host_scene.cpp): initCornellBox()
of RTrace/scene.swift:14-62 converted as RTrace/renderer.swift:49-57 does, the
config-4 sphere scene, and the seeded per-pixel Halton offsets that replace
``arc4random() % 2^20`` (renderer.swift:96-101)."""
import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _lib as L
from ._lib import check, lib


@dataclass
class SceneBytes:
    camera: np.ndarray      # 64 B   CameraGPU
    materials: np.ndarray   # 48*T   MaterialGPU[T]
    lights: np.ndarray      # 64*L   SquareLightGPU[L]
    vertices: np.ndarray    # 48*T   simd_float3[3T]
    spheres: np.ndarray     # 80*S   SphereGPU[S]
    num_triangles: int
    num_spheres: int
    num_lights: int


def _make(cap_t, cap_s, cap_l, fill):
    cam = np.zeros(64, np.uint8)
    mats = np.zeros(48 * cap_t, np.uint8)
    verts = np.zeros(48 * cap_t, np.uint8)
    lights = np.zeros(64 * cap_l, np.uint8)
    sph = np.zeros(80 * max(cap_s, 1), np.uint8)
    hb = L.RTHostSceneBytes(cam.ctypes.data, mats.ctypes.data, cap_t, verts.ctypes.data, lights.ctypes.data, cap_l,
                            sph.ctypes.data, cap_s, 0, 0, 0)
    check(fill(C.byref(hb)), "host scene")
    t, s, l = hb.numTriangles, hb.numSpheres, hb.numLights
    return SceneBytes(cam, mats[:48 * t].copy(), lights[:64 * l].copy(), verts[:48 * t].copy(), sph[:80 * s].copy(), t, s, l)


def init_cornell_box(width=800, height=600, light_size=1.0):
    return _make(64, 0, 4, lambda hb: lib.rtHostInitCornellBox(width, height, C.c_float(light_size), hb))


def init_sphere_scene(width, height, seed=12345, num_spheres=256):
    return _make(64, num_spheres, 8, lambda hb: lib.rtHostInitSphereScene(width, height, seed, num_spheres, hb))


def fill_offsets(width, row0, rows, seed=1):
    off = np.empty(width * rows, np.uint32)
    lib.rtHostFillOffsets(off.ctypes.data_as(C.c_void_p), width, row0, rows, seed)
    return off
