"""The compacting kernels (csrc/wave.cuh) find a ray's spheres by walking a uniform 8 x 8 x 8 grid (csrc/sgrid.cuh) instead
of testing all of them.  oracle/rt_sgrid.c restates the grid's build and the walk on the CPU and repeats the sphere part
of EVERY ray query of a frame through it: the answer (hit, primitive id, distance bits) must be the full loop's.
Runs without a GPU; the GPU side compares the grid kernels with the per-pixel kernel, which never uses the grid
(test_gpu_parity.py)."""
import numpy as np

import oracle_lib as O


def test_grid_never_changes_a_query_sphere_frame():
    W, H, spp = 480, 270, 4
    sc = O.spheres(W, H)
    r = O.sgridcheck(sc, W, H, spp)
    ref, _ = O.pathtrace(sc, W, H, spp)
    assert np.array_equal(r["image"], ref)                 # the hook does not disturb the oracle
    assert r["ok"] and r["queries"] > 5_000_000 and r["mismatches"] == 0
    # what the grid is for: a handful of sphere tests per ray instead of 256
    assert r["tests"] < 4 * r["queries"] and r["cells"] < 4 * r["queries"]


def test_ray_that_starts_on_a_cell_boundary():
    """Row 680 of the 1080p frame holds a bounce ray whose origin lies on a cell face: the first step's parameter comes out
    as -8.7e-8, which an earlier form of the walk read as "left the grid" (one wrong pixel in 2 M, found on the GPU)."""
    W, H = 1920, 1080
    sc = O.spheres(W, H)
    r = O.sgridcheck(sc, W, H, 4, y0=680, y1=681)
    assert r["ok"] and r["queries"] > 90_000 and r["mismatches"] == 0


def test_grid_other_sphere_counts_and_overflow():
    """16 and 64 spheres walk the grid too; a scene that puts more than eight spheres into one cell is flagged, and the
    kernels then trace it by brute force (ok = False: the checker has nothing to repeat)."""
    W, H = 160, 90
    for n in (16, 64):
        sc = O.spheres(W, H, n=n)
        r = O.sgridcheck(sc, W, H, 4)
        assert r["ok"] and r["mismatches"] == 0 and r["queries"] > 100_000
    sc = O.spheres(W, H, n=64)
    sph = sc["spheres"].view(np.float32).reshape(-1, 20)
    sph[:12, 0:3] = sph[0, 0:3]                             # twelve spheres on one spot
    r = O.sgridcheck(sc, W, H, 1)
    assert not r["ok"] and r["queries"] == 0
