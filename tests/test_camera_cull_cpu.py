"""The compacting kernels test a warp's camera rays only against the primitives that camera_candidates
(gpuraytracer_b200/csrc/pathtrace.cu) keeps for the warp's 8x4 pixel tile: a triangle whose projection stays more than
one pixel clear of the tile is dropped unless it is seen nearly edge-on, reaches behind the camera or is a speck.
This file restates the triangle part of that rule in numpy fp32 and checks its geometry without a GPU: for random
cameras (inside and outside the room, any direction, narrow and wide) and random tiles and jitters, the triangle the
oracle's brute-force loop hits is never one the rule drops.  (The device's own lists are checked through whole images
by tests/test_gpu_parity.py::test_camera_candidate_lists_moved_cameras.)"""
import ctypes as C

import numpy as np

import oracle_lib as O

f32 = np.float32


def _norm(a):
    return (a / f32(np.sqrt(f32(a @ a)))).astype(f32)


def camera_frame(cam):
    """generateCameraRay's per-launch part, sampling.metal:131-139 (aspect = integer division, :132)"""
    c = cam.view(f32); ci = cam.view(np.int32)
    aspect = f32(ci[12] // ci[13])
    half_w = f32(np.tan(c[14] / f32(2.0)))
    w = (-_norm(c[4:7])).astype(f32)
    u = _norm(np.cross(c[8:11], w).astype(f32))
    v = _norm(np.cross(w, u).astype(f32))
    return dict(o=c[0:3].copy(), u=u, v=v, w=w, half_w=half_w, half_h=f32(half_w / aspect), res_x=f32(ci[12]), res_y=f32(ci[13]))


def kept(cf, tri, x0, x1, y0, y1):
    """camera_candidates, triangle part: tri = (v0, v1, v2); the tile covers pixels x0..x1, y0..y1"""
    q = (tri - cf["o"]).astype(f32)                       # q0 = v0 - o, q1 = v1 - o, q2 = v2 - o
    n = np.cross(tri[2] - tri[0], tri[1] - tri[0]).astype(f32)           # n = e2 x e1
    zc = -(q @ cf["w"]); xc = q @ cf["u"]; yc = q @ cf["v"]
    with np.errstate(all="ignore"):
        X = (xc / (zc * cf["half_w"]) + f32(1.0)) * (f32(0.5) * cf["res_x"])
        Y = (f32(1.0) - yc / (zc * cf["half_h"])) * (f32(0.5) * cf["res_y"])
    far = np.sqrt(max(q[0] @ q[0], q[1] @ q[1], q[2] @ q[2]))
    safe = zc.min() > 0.01 * far and abs(q[0] @ n) > 0.01 * far * np.sqrt(n @ n)
    outside = X.max() < x0 - 1.0 or X.min() > x1 + 2.0 or Y.max() < y0 - 1.0 or Y.min() > y1 + 2.0
    return not (safe and outside)


def test_the_rule_never_drops_the_triangle_a_camera_ray_hits():
    W, H = 160, 120
    g = np.random.default_rng(21)
    o = np.zeros(3, f32); d = np.zeros(3, f32)
    dist = C.c_float(); pid = C.c_uint32()
    checked = dropped_some = 0
    for scene in range(60):
        sc = O.cornell(W, H)
        cam = sc["camera"]; c = cam.view(f32)
        inside = scene % 3 != 0
        c[0:3] = g.uniform(-2.3, 2.3, 3) if inside else g.uniform(-1.0, 1.0, 3) + (0, 0, g.uniform(4.0, 12.0))
        c[4:7] = g.normal(size=3) if inside else g.uniform(-0.3, 0.3, 3) + (0, 0, -1.0)
        up = g.normal(size=3) if scene % 4 == 0 else np.array([0.0, 1.0, 0.0])
        if abs(np.dot(up, c[4:7])) > 0.95 * np.linalg.norm(up) * np.linalg.norm(c[4:7]):
            up = np.array([0.3, 1.0, 0.1])
        c[8:11] = up
        c[14] = g.uniform(0.2, 2.4)                        # horizontal field of view, radians
        cf = camera_frame(cam)
        v = sc["vertices"].view(f32).reshape(36, 3, 4)[:, :, :3]
        for tile in range(40):
            x0 = int(g.integers(0, W // 8)) * 8; y0 = int(g.integers(0, H // 4)) * 4
            keep = [kept(cf, v[t], x0, x0 + 7, y0, y0 + 3) for t in range(36)]
            dropped_some += not all(keep)
            for ray in range(12):
                # corners and edges of the tile as well as its inside: jitter 0 and the largest float below 1
                jx, jy = (f32(g.choice((0.0, np.nextafter(f32(1), f32(0))))), f32(g.random())) if ray < 4 else (f32(g.random()), f32(g.random()))
                px = x0 + (int(g.choice((0, 7))) if ray < 6 else int(g.integers(8)))
                py = y0 + (int(g.choice((0, 3))) if ray < 6 else int(g.integers(4)))
                O.lib.rt_oracle_camera_ray(O.vp(cam), px, py, C.c_float(jx), C.c_float(jy), O.vp(o), O.vp(d))
                if O.lib.rt_oracle_intersect(O.vp(sc["vertices"]), 36, None, 0, O.vp(o), O.vp(d), C.c_float(1e-3),
                                             C.c_float(1000.0), 0, C.byref(dist), C.byref(pid)):
                    assert keep[pid.value], (scene, tile, px, py, jx, jy, pid.value)
                    checked += 1
    assert checked > 15_000 and dropped_some > 1_000       # the rule does drop things: the check is not vacuous


def kept_sphere(cf, centre, r, x0, x1, y0, y1):
    """camera_candidates, sphere part: the box bound of x/z and y/z over the sphere's bounding cube"""
    q = (centre - cf["o"]).astype(f32)
    zc = -(q @ cf["w"]); xc = q @ cf["u"]; yc = q @ cf["v"]
    far = np.sqrt(q @ q)
    safe = zc - r > 0.01 * (far + r) and r > 1e-3 * far
    zn, zf = zc - r, zc + r
    with np.errstate(all="ignore"):
        txh = (xc + r) / (zn if xc + r >= 0 else zf); txl = (xc - r) / (zn if xc - r <= 0 else zf)
        tyh = (yc + r) / (zn if yc + r >= 0 else zf); tyl = (yc - r) / (zn if yc - r <= 0 else zf)
        Xh = (txh / cf["half_w"] + 1.0) * (0.5 * cf["res_x"]); Xl = (txl / cf["half_w"] + 1.0) * (0.5 * cf["res_x"])
        Yl = (1.0 - tyh / cf["half_h"]) * (0.5 * cf["res_y"]); Yh = (1.0 - tyl / cf["half_h"]) * (0.5 * cf["res_y"])
    outside = Xh < x0 - 1.0 or Xl > x1 + 2.0 or Yh < y0 - 1.0 or Yl > y1 + 2.0
    return not (safe and outside)


def test_the_rule_never_drops_the_sphere_a_camera_ray_hits():
    """Config 4's scene (256 spheres, 8 light triangles), cameras between the spheres and outside the room."""
    W, H = 160, 120
    g = np.random.default_rng(22)
    o = np.zeros(3, f32); d = np.zeros(3, f32)
    dist = C.c_float(); pid = C.c_uint32()
    checked = dropped = 0
    for scene in range(16):
        sc = O.spheres(W, H)
        T, S = sc["T"], sc["S"]
        cam = sc["camera"]; c = cam.view(f32)
        if scene % 2:
            c[0:3] = g.uniform(-2.2, 2.2, 3); c[4:7] = g.normal(size=3)
        else:
            c[0:3] = g.uniform(-1.0, 1.0, 3) + (0, 0, g.uniform(4.0, 10.0)); c[4:7] = g.uniform(-0.3, 0.3, 3) + (0, 0, -1.0)
        c[14] = g.uniform(0.3, 2.0)
        cf = camera_frame(cam)
        sp = sc["spheres"].view(f32).reshape(S, 20)
        v = sc["vertices"].view(f32).reshape(T, 3, 4)[:, :, :3]
        for tile in range(20):
            x0 = int(g.integers(0, W // 8)) * 8; y0 = int(g.integers(0, H // 4)) * 4
            keep = [kept(cf, v[t], x0, x0 + 7, y0, y0 + 3) for t in range(T)] + \
                   [kept_sphere(cf, sp[s, 0:3], abs(sp[s, 16]), x0, x0 + 7, y0, y0 + 3) for s in range(S)]
            dropped += len(keep) - sum(keep)
            for ray in range(16):
                jx, jy = f32(g.random()), f32(g.random())
                px = x0 + (int(g.choice((0, 7))) if ray < 8 else int(g.integers(8)))
                py = y0 + (int(g.choice((0, 3))) if ray < 8 else int(g.integers(4)))
                O.lib.rt_oracle_camera_ray(O.vp(cam), px, py, C.c_float(jx), C.c_float(jy), O.vp(o), O.vp(d))
                if O.lib.rt_oracle_intersect(O.vp(sc["vertices"]), T, O.vp(sc["spheres"]), S, O.vp(o), O.vp(d), C.c_float(1e-3),
                                             C.c_float(1000.0), 0, C.byref(dist), C.byref(pid)):
                    assert keep[pid.value], (scene, tile, px, py, pid.value)
                    checked += 1
    assert checked > 2_000 and dropped > 40_000
