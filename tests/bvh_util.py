"""Test helper: a finer triangulation of a scene (every triangle cut into n*n), for the BVH kernel."""
import numpy as np


def tessellate(vertices, materials, T, n):
    """vertices: uint8[48*T] (3 x simd_float3 per triangle), materials: uint8[48*T].  Returns (vertices, materials, T*n*n)."""
    v = vertices.view(np.float32).reshape(T, 3, 4)
    m = materials.reshape(T, 48)
    out_v = np.zeros((T * n * n, 3, 4), np.float32)
    out_m = np.zeros((T * n * n, 48), np.uint8)
    k = 0
    inv = np.float32(1.0) / np.float32(n)
    for t in range(T):
        v0, e1, e2 = v[t, 0, :3], v[t, 1, :3] - v[t, 0, :3], v[t, 2, :3] - v[t, 0, :3]

        def pt(i, j):
            return (v0 + (np.float32(i) * inv) * e1 + (np.float32(j) * inv) * e2).astype(np.float32)
        for i in range(n):
            for j in range(n - i):
                tris = [(pt(i, j), pt(i + 1, j), pt(i, j + 1))]
                if i + j + 2 <= n:
                    tris.append((pt(i + 1, j), pt(i + 1, j + 1), pt(i, j + 1)))
                for a, b, c in tris:
                    out_v[k, 0, :3] = a; out_v[k, 1, :3] = b; out_v[k, 2, :3] = c
                    out_m[k] = m[t]
                    k += 1
    assert k == T * n * n
    return out_v.reshape(-1).view(np.uint8).copy(), out_m.reshape(-1).copy(), k


LEAF = 0x80000000


def check_nodes(nodes, order, verts, T, max_leaf=12):
    """Walks a hierarchy in the layout of csrc/host_bvh.h (nodes: uint32 [n, 16]) from inner node 0: every triangle is in
    exactly one leaf, leaves hold at most max_leaf, and every box on the way down to a triangle contains it.
    Returns (depth, leaves, mean triangles per leaf)."""
    assert sorted(np.asarray(order).tolist()) == list(range(T))
    v = verts.view(np.float32).reshape(T, 3, 4)[:, :, :3]
    f = nodes.view(np.float32)
    seen = np.zeros(T, bool)
    stack = [(0, np.full(3, -np.inf, np.float32), np.full(3, np.inf, np.float32), 1)]
    maxd, leaves = 0, 0
    while stack:
        ref, lo, hi, d = stack.pop()
        maxd = max(maxd, d)
        if ref & LEAF:
            first, count = ref & 0x07ffffff, ((ref >> 27) & 15) + 1
            assert count <= max_leaf
            leaves += 1
            ids = np.asarray(order[first:first + count])
            assert not seen[ids].any()
            seen[ids] = True
            tri = v[ids]
            assert np.all(tri >= lo) and np.all(tri <= hi), (first, count)
            continue
        nd, fd = nodes[ref], f[ref]
        for lo_c, hi_c, r in ((fd[0:3], fd[4:7], int(nd[3])), (fd[8:11], fd[12:15], int(nd[7]))):
            if np.all(lo_c <= hi_c):                       # (an empty box marks the unused child of a one-leaf tree)
                stack.append((r, np.maximum(lo, lo_c), np.minimum(hi, hi_c), d + 1))
    assert seen.all()
    return maxd, leaves, T / max(leaves, 1)


def tessellate_fast(vertices, materials, T, n):
    """tessellate() vectorised per source triangle (same points: v0 + (i/n) e1 + (j/n) e2 in fp32; the triangles come in
    another order), for scenes of a million triangles."""
    v = vertices.view(np.float32).reshape(T, 3, 4)
    m = materials.reshape(T, 48)
    inv = np.float32(1.0) / np.float32(n)
    out_v, out_m = [], []
    for t in range(T):
        v0, e1, e2 = v[t, 0, :3], v[t, 1, :3] - v[t, 0, :3], v[t, 2, :3] - v[t, 0, :3]

        def pt(i, j):
            return (v0 + (i.astype(np.float32) * inv)[:, None] * e1 + (j.astype(np.float32) * inv)[:, None] * e2).astype(np.float32)
        i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
        up = (i + j) < n                                            # (i, j), (i+1, j), (i, j+1)
        dn = (i + j + 2) <= n                                       # (i+1, j), (i+1, j+1), (i, j+1)
        iu, ju, idn, jdn = i[up], j[up], i[dn], j[dn]
        tri = np.zeros((iu.size + idn.size, 3, 4), np.float32)
        tri[:iu.size, 0, :3], tri[:iu.size, 1, :3], tri[:iu.size, 2, :3] = pt(iu, ju), pt(iu + 1, ju), pt(iu, ju + 1)
        tri[iu.size:, 0, :3], tri[iu.size:, 1, :3], tri[iu.size:, 2, :3] = pt(idn + 1, jdn), pt(idn + 1, jdn + 1), pt(idn, jdn + 1)
        out_v.append(tri)
        out_m.append(np.tile(m[t], (tri.shape[0], 1)))
    V, M = np.concatenate(out_v), np.concatenate(out_m)
    assert V.shape[0] == T * n * n
    return V.reshape(-1).view(np.uint8).copy(), M.reshape(-1).copy(), V.shape[0]
