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


def random_sphere_scene(W, H, seed, n):
    """Config 4's room and lights with `n` spheres thrown at random: centres anywhere in the room (some poke through a wall),
    radii log-uniform from 0.02 to 0.9 (a few span several cells of the grid, some cells overflow for large n), every tenth
    sphere a copy of its neighbour's centre (equal distances: the tie rule)."""
    sc = O.spheres(W, H, n=n)
    g = np.random.default_rng(seed)
    s = sc["spheres"].view(np.float32).reshape(-1, 20)
    s[:, 0:3] = g.uniform(-2.4, 2.4, (n, 3)).astype(np.float32)
    s[:, 16] = np.exp(g.uniform(np.log(0.02), np.log(0.9), n)).astype(np.float32)
    for j in range(9, n, 10):
        s[j, 0:3] = s[j - 1, 0:3]
    return sc


def test_grid_never_changes_a_query_random_sphere_scenes():
    W, H = 120, 68
    usable = total = 0
    for seed, n in ((1, 1), (2, 7), (3, 40), (4, 40), (5, 120), (6, 256), (7, 256), (8, 19)):
        sc = random_sphere_scene(W, H, seed, n)
        r = O.sgridcheck(sc, W, H, 4, bounces=4)
        assert r["mismatches"] == 0, (seed, n, r["mismatches"], r["first_bad"])
        usable += r["ok"]; total += r["queries"]
    assert usable >= 5 and total > 500_000          # the crowded ones overflow a cell and are traced by the full loop instead
