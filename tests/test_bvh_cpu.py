"""The host-side BVH builder (csrc/host_bvh.cpp) without a GPU: every triangle sits in exactly one
leaf, leaves hold at most twelve, and every box on the way from the root to a triangle contains it (grown by the margin)."""
import ctypes as C

import numpy as np

import oracle_lib as O
from bvh_util import tessellate
from gpuraytracer_b200 import _lib as L

LEAF = 0x80000000


def build(verts, T):
    nodes = np.zeros((max(T, 2), 16), np.uint32)
    order = np.zeros(T, np.uint32)
    n = C.c_uint32(); margin = C.c_float()
    L.lib.rtDebugBuildBvh.restype = C.c_int
    depth = L.lib.rtDebugBuildBvh(verts.ctypes.data_as(C.c_void_p), C.c_uint32(T), nodes.ctypes.data_as(C.c_void_p),
                                  C.c_uint32(nodes.shape[0]), order.ctypes.data_as(C.c_void_p), C.byref(n), C.byref(margin))
    assert depth > 0
    return nodes[:n.value], order, depth, margin.value


def check_tree(verts, T):
    nodes, order, depth, margin = build(verts, T)
    assert sorted(order.tolist()) == list(range(T))
    v = verts.view(np.float32).reshape(T, 3, 4)[:, :, :3]
    f = nodes.view(np.float32)
    seen = np.zeros(T, bool)
    stack = [(0, np.full(3, -np.inf, np.float32), np.full(3, np.inf, np.float32), 1)]
    maxd = 0
    while stack:
        ref, lo, hi, d = stack.pop()
        maxd = max(maxd, d)
        if ref & LEAF:
            first, count = ref & 0x07ffffff, ((ref >> 27) & 15) + 1
            assert count <= 12
            for s in range(first, first + count):
                tri = v[order[s]]
                assert not seen[order[s]]; seen[order[s]] = True
                assert np.all(tri >= lo) and np.all(tri <= hi)
            continue
        nd = nodes[ref]; fd = f[ref]
        for lo_c, hi_c, r in ((fd[0:3], fd[4:7], int(nd[3])), (fd[8:11], fd[12:15], int(nd[7]))):
            if np.all(lo_c <= hi_c):                       # (an empty box marks the unused child of a one-leaf tree)
                stack.append((r, np.maximum(lo, lo_c - 0), np.minimum(hi, hi_c + 0), d + 1))
    assert seen.all() or T == 0
    assert maxd <= depth + 1 and depth <= 44
    assert margin > 0
    return len(nodes), depth


def test_builder_on_cornell_and_tessellations():
    sc = O.cornell(64, 64)
    assert check_tree(sc["vertices"], 36)[0] >= 3
    for n in (3, 6):
        tv, tm, T = tessellate(sc["vertices"], sc["materials"], 36, n)
        nn, depth = check_tree(tv, T)
        assert nn <= T and depth <= 16


def test_builder_small_and_degenerate():
    sc = O.cornell(64, 64)
    for T in (1, 2, 4, 5, 12, 13, 25):
        check_tree(sc["vertices"][:48 * T].copy(), T)
    same = np.tile(sc["vertices"][:48], 37).copy()        # 37 copies of one triangle: all centroids equal
    check_tree(same, 37)


def _paths(nodes):
    """leaf slot -> the (lo, hi) boxes, as stored (grown), on the way from the root to the slot's leaf"""
    f = nodes.view(np.float32)
    out = {}
    stack = [(0, [])]
    while stack:
        ref, boxes = stack.pop()
        if ref & LEAF:
            first, count = ref & 0x07ffffff, ((ref >> 27) & 15) + 1
            for s in range(first, first + count):
                out[s] = np.array(boxes, np.float32)
            continue
        nd, fd = nodes[ref], f[ref]
        for lo_c, hi_c, r in ((fd[0:3], fd[4:7], int(nd[3])), (fd[8:11], fd[12:15], int(nd[7]))):
            if np.all(lo_c <= hi_c):
                stack.append((r, boxes + [(lo_c.copy(), hi_c.copy())]))
    return out


def _slab(boxes, o, d):
    """csrc/bvh.cuh bvh_inv + bvh_box in numpy fp32, for all boxes of a path at once: (kept, entry distance)"""
    f32 = np.float32
    with np.errstate(divide="ignore", over="ignore", invalid="ignore"):
        inv = np.where(np.abs(d) > f32(1e-30), f32(1.0) / d, np.copysign(f32(1e30), d)).astype(f32)
        t0 = ((boxes[:, 0] - o) * inv).astype(f32)
        t1 = ((boxes[:, 1] - o) * inv).astype(f32)
    near, far = np.fmin(t0, t1), np.fmax(t0, t1)
    tn = np.fmax(np.fmax(near[:, 0], near[:, 1]), np.fmax(near[:, 2], f32(0.0)))
    tf = np.fmin(np.fmin(far[:, 0], far[:, 1]), far[:, 2])
    return ~(tn > tf * f32(1.00001) + f32(1e-30)), tn


def test_no_ray_loses_its_hit_to_a_box():
    """Seeded fuzz of the traversal's two skip rules (csrc/bvh.cuh bvh_box / bvh_step, restated in numpy fp32) against
    the oracle's brute-force answer: for rays from inside the room -- random, axis-parallel (zero direction components:
    the 0 * inf case), grazing along walls, and starting ON a wall -- every box on the way from the root to the leaf of
    the triangle the brute-force loop returns is entered (slab test) and entered no later than the hit (the pruning
    limit at its tightest), so the traversal must reach that triangle whatever order it walks in."""
    sc = O.cornell(64, 64)
    g = np.random.default_rng(5)
    dist = C.c_float(); pid = C.c_uint32()
    checked = 0
    for n in (1, 3, 8):
        verts, T = (sc["vertices"], 36) if n == 1 else tessellate(sc["vertices"], sc["materials"], 36, n)[::2]
        nodes, order, depth, margin = build(verts, T)
        paths = _paths(nodes)
        slot_of = np.empty(T, np.int64); slot_of[order] = np.arange(T)
        for k in range(1500):
            o = g.uniform(-2.4, 2.4, 3).astype(np.float32)
            d = g.normal(size=3)
            kind = k % 5
            if kind == 1:                                   # axis-parallel
                d = np.zeros(3); d[g.integers(3)] = g.choice((-1.0, 1.0))
            elif kind == 2:                                 # in a coordinate plane
                d[g.integers(3)] = 0.0
            elif kind == 3:                                 # grazing: almost parallel to a wall, starting close to it
                a = g.integers(3); d[a] *= 1e-4; o[a] = np.float32(g.choice((-2.5, 2.5)) * (1 - 1e-3))
            elif kind == 4:                                 # origin exactly in a wall's plane (shared box faces)
                o[g.integers(3)] = np.float32(g.choice((-2.5, 2.5)))
            d = (d / np.linalg.norm(d)).astype(np.float32)
            hit = O.lib.rt_oracle_intersect(O.vp(verts), T, None, 0, O.vp(o), O.vp(d), C.c_float(1e-3), C.c_float(1000.0),
                                            0, C.byref(dist), C.byref(pid))
            if not hit:
                continue
            kept, tn = _slab(paths[int(slot_of[pid.value])], o, d)
            assert kept.all(), (n, k, o, d, pid.value)
            assert np.all(tn <= np.float32(dist.value) * np.float32(1.00005)), (n, k, o, d, pid.value, tn.max(), dist.value)
            checked += 1
    assert checked > 3000


def test_closest_hit_does_not_depend_on_the_order_triangles_are_met_in():
    """The BVH kernel meets the triangles in tree order and picks the closest by "smaller cross-multiplied distance, on
    an exact tie the lower id" (oracle hook rt_oracle_intersect_ordered = csrc/bvh.cuh bvh_leaf_closest).  That must be
    the ascending strict-less loop's answer (rt_oracle_intersect) whatever the order -- checked on rays made to tie:
    axis-parallel rays through the grid vertices and edges of tessellated walls (up to six triangles share the hit
    point, distance bit-equal), rays through vertices from random origins, and plain random rays."""
    sc = O.cornell(64, 64)
    g = np.random.default_rng(9)
    O.lib.rt_oracle_intersect_ordered.restype = C.c_int
    d0, p0, d1, p1 = C.c_float(), C.c_uint32(), C.c_float(), C.c_uint32()
    ties = 0
    for n in (1, 4, 8):
        verts, T = (sc["vertices"], 36) if n == 1 else tessellate(sc["vertices"], sc["materials"], 36, n)[::2]
        v = verts.view(np.float32).reshape(T, 3, 4)[:, :, :3]
        for k in range(400):
            t = int(g.integers(T)); corner = v[t, int(g.integers(3))]
            kind = k % 4
            if kind == 0:                                   # axis-parallel through a vertex
                a = int(g.integers(3)); o = corner.copy(); o[a] = np.float32(g.uniform(-2.0, 2.0))
                d = np.zeros(3, np.float32); d[a] = np.sign(corner[a] - o[a]) or 1.0
            elif kind == 1:                                 # axis-parallel through an edge midpoint
                mid = ((corner + v[t, int(g.integers(3))]) * np.float32(0.5)).astype(np.float32)
                a = int(g.integers(3)); o = mid.copy(); o[a] = np.float32(g.uniform(-2.0, 2.0))
                d = np.zeros(3, np.float32); d[a] = np.sign(mid[a] - o[a]) or 1.0
            elif kind == 2:                                 # towards a vertex from anywhere
                o = g.uniform(-2.4, 2.4, 3).astype(np.float32)
                d = (corner - o).astype(np.float32); d = (d / np.float32(np.linalg.norm(d))).astype(np.float32)
            else:
                o = g.uniform(-2.4, 2.4, 3).astype(np.float32)
                d = g.normal(size=3); d = (d / np.linalg.norm(d)).astype(np.float32)
            hit = O.lib.rt_oracle_intersect(O.vp(verts), T, None, 0, O.vp(o), O.vp(d), C.c_float(1e-3), C.c_float(1000.0),
                                            0, C.byref(d0), C.byref(p0))
            for perm in range(6):                           # ascending, descending, four random orders
                ids = (np.arange(T) if perm == 0 else np.arange(T)[::-1] if perm == 1 else g.permutation(T)).astype(np.uint32)
                h = O.lib.rt_oracle_intersect_ordered(O.vp(verts), O.vp(np.ascontiguousarray(ids)), T, O.vp(o), O.vp(d),
                                                      C.c_float(1e-3), C.c_float(1000.0), C.byref(d1), C.byref(p1))
                assert h == hit, (n, k, perm)
                if hit:
                    assert p1.value == p0.value and np.float32(d1.value).view(np.uint32) == np.float32(d0.value).view(np.uint32), \
                        (n, k, perm, p0.value, p1.value, d0.value, d1.value)
            if hit and kind < 2:
                # was there a tie to break?  (the next triangle alone gives the same distance: the other half of a quad)
                one = np.array([p0.value + 1], np.uint32)
                if p0.value + 1 < T and O.lib.rt_oracle_intersect_ordered(O.vp(verts), O.vp(one), 1, O.vp(o), O.vp(d), C.c_float(1e-3),
                                                                         C.c_float(1000.0), C.byref(d1), C.byref(p1)):
                    ties += int(np.float32(d1.value) == np.float32(d0.value))
    assert ties > 50
