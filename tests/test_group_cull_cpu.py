"""The sorted kernel (csrc/sorted.cuh) tests a ray only against the quads of the object groups whose grown box the
ray's line does not stay clear of.  oracle/rt_groups.c restates the grouping and the box rule on the CPU and repeats
EVERY ray query of a frame against the reduced quad list: the answer (hit, primitive id, distance bits) must be
the full loop's.  Runs without a GPU; the GPU side compares the device's groups with these (test_gpu_parity.py)."""
import numpy as np

import oracle_lib as O


def test_cornell_groups_are_room_boxes_light():
    g = O.groups(O.cornell(64, 64))
    K, always, ok = int(g[0]), int(g[1]), int(g[2])
    assert (K, ok) == (2, 1)
    assert always == 0b100000000000011111                  # quads 0-4 (room) and 17 (light): always tested
    assert int(g[4 + 3]) == 0b000000011111100000           # tall box: quads 5-10
    assert int(g[12 + 3]) == 0b011111100000000000          # short box: quads 11-16
    f = g.view(np.float32)
    # tall box: centre (-1, -1.15, -1.5), 1.2 x 2.8 x 1.2 turned by 75 degrees, grown by 5.05 / 128 (the boxes reach 0.05 below the floor)
    assert np.allclose(f[4:7], [-1.0, -1.15, -1.5], atol=1e-6)
    assert np.allclose(f[8:11], [0.6 * (np.cos(np.pi / 2.4) + np.sin(np.pi / 2.4)) + 5.05 / 128, 1.4 + 5.05 / 128,
                                 0.6 * (np.cos(np.pi / 2.4) + np.sin(np.pi / 2.4)) + 5.05 / 128], atol=1e-5)


def test_culling_never_changes_a_query_cornell_frame():
    W, H, spp = 960, 540, 8
    sc = O.cornell(W, H)
    r = O.cullcheck(sc, W, H, spp)
    ref, _ = O.pathtrace(sc, W, H, spp)
    assert np.array_equal(r["image"], ref)                 # the hook does not disturb the oracle
    assert r["queries"] > 15_000_000 and r["mismatches"] == 0
    # what the sorting is for: two thirds of the bounce steps need neither box
    dual = r["keys_dual"] / r["keys_dual"].sum()
    assert dual[0] > 0.6 and dual[3] < 0.03
    assert r["culled_tests"] < 0.55 * r["full_tests"]


def test_culling_never_changes_a_query_other_scenes():
    """Free-floating and clustered quads in place of the boxes (more groups than key bits), a camera inside the room."""
    W, H, spp = 160, 120, 6
    for mode in ("floating", "clusters"):
        sc = O.cornell(W, H)
        f = sc["vertices"].view(np.float32)
        g = np.random.default_rng(3)

        def quad(t, a, e1, e2):
            for k, p in enumerate((a, a + e1, a + e1 + e2, a, a + e1 + e2, a + e2)):
                f[(3 * t + k) * 4:(3 * t + k) * 4 + 3] = p
        a = None
        for q in range(12):
            if mode == "floating" or q % 4 == 0:
                a = g.uniform(-2.0, 2.0, 3).astype(np.float32)
            quad(10 + 2 * q, a, g.uniform(-0.9, 0.9, 3).astype(np.float32), g.uniform(-0.9, 0.9, 3).astype(np.float32))
        cam = sc["camera"].view(np.float32)
        cam[0:3] = (0.3, -0.4, 2.0); cam[4:7] = (0.1, -0.1, -1.0); cam[14] = 1.4
        grp = O.groups(sc)
        assert int(grp[2]) == 1 and int(grp[0]) == (4 if mode == "floating" else 3)
        r = O.cullcheck(sc, W, H, spp, bounces=4)
        assert r["queries"] > 300_000 and r["mismatches"] == 0, (mode, r["mismatches"])
