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
