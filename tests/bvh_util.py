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
