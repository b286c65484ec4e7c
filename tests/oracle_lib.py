"""ctypes access to oracle/liboracle.so -- the CHECKER.  Only tests/, smoke() and
bench.py's cpu_baseline / --impl reference legs import this module."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_PATH = os.environ.get("ORACLE_LIB") or os.path.join(ROOT, "oracle", "liboracle.so")   # (override: the sanitizer build, oracle/Makefile)
if not os.path.exists(_PATH):
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")])
lib = C.CDLL(_PATH)


class RTLaunchParams(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("row0", C.c_int32), ("rows", C.c_int32),
                ("spp", C.c_uint32), ("firstSample", C.c_uint32), ("bounces", C.c_int32),
                ("accumulate", C.c_int32), ("numTriangles", C.c_uint32), ("numSpheres", C.c_uint32),
                ("numLights", C.c_uint32), ("cameraRays", C.c_uint32), ("misSamples", C.c_uint32),
                ("variant", C.c_uint32), ("tileRows", C.c_uint32), ("tileStride", C.c_uint32)]


lib.rt_oracle_halton.restype = C.c_float
lib.rt_oracle_halton.argtypes = [C.c_uint32, C.c_uint32]
lib.rt_oracle_hash.restype = C.c_uint32
lib.rt_oracle_hash.argtypes = [C.c_uint32]
lib.rt_oracle_max_threads.restype = C.c_int
lib.rt_oracle_set_light_normals.restype = None
lib.rt_oracle_set_light_normals.argtypes = [C.c_void_p]


def vp(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def cornell(width, height, light_size=1.0):
    cam = np.zeros(64, np.uint8); mats = np.zeros(36 * 48, np.uint8)
    light = np.zeros(64, np.uint8); verts = np.zeros(108 * 16, np.uint8)
    lib.rt_fixture_cornell(width, height, C.c_float(light_size), vp(cam), vp(mats), vp(light), vp(verts))
    return dict(camera=cam, materials=mats, lights=light, vertices=verts, spheres=None, T=36, S=0, L=1)


def spheres(width, height, seed=12345, n=256):
    cam = np.zeros(64, np.uint8); mats = np.zeros(18 * 48, np.uint8)
    lights = np.zeros(4 * 64, np.uint8); verts = np.zeros(54 * 16, np.uint8); sph = np.zeros(n * 80, np.uint8)
    lib.rt_fixture_spheres(width, height, C.c_uint32(seed), C.c_uint32(n), vp(cam), vp(mats), vp(lights), vp(verts), vp(sph))
    return dict(camera=cam, materials=mats, lights=lights, vertices=verts, spheres=sph, T=18, S=n, L=4)


def offsets(width, row0, rows, seed=1):
    off = np.zeros(width * rows, np.uint32)
    lib.rt_fixture_offsets(vp(off), width, row0, rows, C.c_uint32(seed))
    return off


MIS, MIS_LIGHT_ONLY, MIS_COSINE_ONLY = 0x400, 0x400 | 0x800, 0x400 | 0x1000      # RT_VARIANT_MIS; the checker-only selectors


def pathtrace(scene, width, height, spp, *, row0=0, rows=None, first_sample=0, bounces=3, accumulate=False,
              out=None, off=None, seed=1, nthreads=0, tile_rows=0, tile_stride=0, light_extent=False, mis=0,
              light_normals=None):
    """Returns (image float32 [rows,width,4], counters uint64[8]).  light_extent: RT_VARIANT_LIGHT_EXTENT (0x200);
    mis: 0, MIS, MIS_LIGHT_ONLY or MIS_COSINE_ONLY; light_normals: float32 [L, 4] or None (the reference's (0,-1,0))."""
    rows = height - row0 if rows is None else rows
    if off is None:
        off = offsets(width, row0, rows, seed)
    if out is None:
        out = np.zeros((rows, width, 4), np.float32)
    cnt = np.zeros(8, np.uint64)
    p = RTLaunchParams(width, height, row0, rows, spp, first_sample, bounces, int(accumulate),
                       scene["T"], scene["S"], scene["L"], 6, 300, (0x200 if light_extent else 0) | int(mis), tile_rows, tile_stride)
    normals = None if light_normals is None else np.ascontiguousarray(light_normals, np.float32).reshape(-1, 4)
    lib.rt_oracle_set_light_normals(vp(normals))
    try:
        rc = lib.rt_oracle_pathtrace(vp(scene["camera"]), vp(scene["materials"]), vp(scene["lights"]),
                                     vp(scene["vertices"]), vp(scene["spheres"]), vp(off), vp(out), C.byref(p),
                                     vp(cnt), nthreads)
    finally:
        lib.rt_oracle_set_light_normals(None)
    if rc != 0:
        raise RuntimeError(f"rt_oracle_pathtrace returned {rc}")
    return out, cnt


def drawtriangle(scene, width, height, *, row0=0, rows=None, camera_rays=6, mis_samples=300, nthreads=0,
                 tile_rows=0, tile_stride=0):
    """Mode B (drawTriangle, MIS).  Returns (pixels uint8 [rows,width,4], text float32 [rows,width,3], counters)."""
    rows = height - row0 if rows is None else rows
    pixels = np.zeros((rows, width, 4), np.uint8)
    text = np.zeros((rows, width, 3), np.float32)
    cnt = np.zeros(8, np.uint64)
    p = RTLaunchParams(width, height, row0, rows, 0, 0, 0, 0, scene["T"], 0, scene["L"], camera_rays, mis_samples,
                       0, tile_rows, tile_stride)
    rc = lib.rt_oracle_drawtriangle(vp(scene["camera"]), vp(scene["materials"]), vp(scene["lights"]),
                                    vp(scene["vertices"]), vp(pixels), vp(text), C.byref(p), vp(cnt), nthreads)
    if rc != 0:
        raise RuntimeError(f"rt_oracle_drawtriangle returned {rc}")
    return pixels, text, cnt


def flops(cnt, T, S):
    """F = 45*T*(Rc+Rs) + 18*S*(Rc+Rs) + 200*B + 60*P."""
    P, Rc, Rs, B = (int(cnt[i]) for i in range(4))
    return 45 * T * (Rc + Rs) + 18 * S * (Rc + Rs) + 200 * B + 60 * P


def cullcheck_modeb(scene, width, height, *, camera_rays=6, mis_samples=300, nthreads=0):
    """Traces a drawTriangle frame; every closest-hit query is repeated against the quads the group / shell rule leaves.
    Returns dict(queries, mismatches, culled_tests, full_tests, seg_queries, seg_in_shell, text)."""
    pixels = np.zeros((height, width, 4), np.uint8)
    text = np.zeros((height, width, 3), np.float32)
    stats = np.zeros(6, np.uint64)
    p = RTLaunchParams(width, height, 0, height, 0, 0, 0, 0, scene["T"], 0, scene["L"], camera_rays, mis_samples, 0, 0, 0)
    rc = lib.rt_oracle_cullcheck_modeb(vp(scene["camera"]), vp(scene["materials"]), vp(scene["lights"]), vp(scene["vertices"]),
                                       vp(pixels), vp(text), C.byref(p), vp(stats), nthreads)
    if rc != 0:
        raise RuntimeError(f"rt_oracle_cullcheck_modeb returned {rc}")
    return dict(queries=int(stats[0]), mismatches=int(stats[1]), culled_tests=int(stats[2]), full_tests=int(stats[3]),
                seg_queries=int(stats[4]), seg_in_shell=int(stats[5]), text=text, pixels=pixels)


def shellcheck(scene, width, height, spp, *, bounces=3, seed=1, nthreads=0):
    """Traces the frame; every shadow query from inside the shrunk shell box is repeated without the shell triangles
    (oracle/rt_groups.c: the several-lights passes' shell rule).  Returns dict(shadow, skipped, mismatches, groups)."""
    off = offsets(width, 0, height, seed)
    out = np.zeros((height, width, 4), np.float32)
    stats = np.zeros(3, np.uint64)
    p = RTLaunchParams(width, height, 0, height, spp, 0, bounces, 0, scene["T"], scene["S"], scene["L"], 6, 300, 0, 0, 0)
    rc = lib.rt_oracle_shellcheck(vp(scene["camera"]), vp(scene["materials"]), vp(scene["lights"]), vp(scene["vertices"]),
                                  vp(scene["spheres"]), vp(off), vp(out), C.byref(p), vp(stats), nthreads)
    if rc != 0:
        raise RuntimeError(f"rt_oracle_shellcheck returned {rc}")
    grp = np.zeros(52, np.uint32)
    (lib.rt_oracle_groups_spheres if scene["S"] else lib.rt_oracle_groups)(vp(scene["vertices"]), C.c_uint32(scene["T"]), vp(scene["lights"]),
                                                                           C.c_uint32(scene["L"]), vp(grp))
    return dict(shadow=int(stats[0]), skipped=int(stats[1]), mismatches=int(stats[2]), groups=grp, image=out)


def sgridcheck(scene, width, height, spp, *, bounces=3, seed=1, y0=0, y1=None, nthreads=0):
    """Traces rows [y0, y1) of the frame; the sphere part of every ray query is repeated by walking the sphere grid
    (oracle/rt_sgrid.c).  Returns dict(queries, mismatches, cells, tests, ok, first_bad[8], image)."""
    rows = (height if y1 is None else y1) - y0
    off = offsets(width, y0, rows, seed)
    out = np.zeros((rows, width, 4), np.float32)
    stats = np.zeros(5, np.uint64)
    bad = np.zeros(8, np.float32)
    p = RTLaunchParams(width, height, y0, rows, spp, 0, bounces, 0, scene["T"], scene["S"], scene["L"], 6, 300, 0, 0, 0)
    rc = lib.rt_oracle_sgridcheck(vp(scene["camera"]), vp(scene["materials"]), vp(scene["lights"]), vp(scene["vertices"]),
                                  vp(scene["spheres"]), vp(off), vp(out), C.byref(p), vp(stats), vp(bad), nthreads)
    if rc != 0:
        raise RuntimeError(f"rt_oracle_sgridcheck returned {rc}")
    return dict(queries=int(stats[0]), mismatches=int(stats[1]), cells=int(stats[2]), tests=int(stats[3]), ok=bool(stats[4]),
                first_bad=bad, image=out)


def groups(scene):
    """The 52 words of the prepared scene's "group" section, restated on the CPU (oracle/rt_groups.c)."""
    out = np.zeros(52, np.uint32)
    lib.rt_oracle_groups(vp(scene["vertices"]), C.c_uint32(scene["T"]), vp(scene["lights"]), C.c_uint32(scene["L"]), vp(out))
    return out


def cullcheck(scene, width, height, spp, *, bounces=3, seed=1, nthreads=0):
    """Traces the frame; every ray query is repeated against the quads group culling leaves.
    Returns dict(queries, mismatches, culled_tests, full_tests, keys_dual[16], keys_single[16])."""
    off = offsets(width, 0, height, seed)
    out = np.zeros((height, width, 4), np.float32)
    stats = np.zeros(36, np.uint64)
    p = RTLaunchParams(width, height, 0, height, spp, 0, bounces, 0, scene["T"], 0, scene["L"], 6, 300, 0, 0, 0)
    rc = lib.rt_oracle_cullcheck(vp(scene["camera"]), vp(scene["materials"]), vp(scene["lights"]), vp(scene["vertices"]),
                                 vp(off), vp(out), C.byref(p), vp(stats), nthreads)
    if rc != 0:
        raise RuntimeError(f"rt_oracle_cullcheck returned {rc}")
    shell = np.zeros(2, np.uint64)
    lib.rt_oracle_cullcheck_shell(vp(shell))
    return dict(queries=int(stats[0]), mismatches=int(stats[1]), culled_tests=int(stats[2]), full_tests=int(stats[3]),
                shadow_queries=int(shell[0]), shadow_in_shell=int(shell[1]),
                keys_dual=stats[4:20].astype(np.int64), keys_single=stats[20:36].astype(np.int64), image=out)
