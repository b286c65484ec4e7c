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
    assert int(g[3]) == 0b000000000000011111               # the shell: the five walls (the light hangs 0.01 below the ceiling)
    assert np.allclose(g.view(np.float32)[36:39], 0.0) and np.allclose(g.view(np.float32)[40:43], 2.5 - 5.05 / 16384, atol=1e-6)
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
    assert r["culled_tests"] < 0.42 * r["full_tests"]
    # the shell rule: practically every shadow ray of the box starts and ends inside the room
    assert r["shadow_in_shell"] > 0.999 * r["shadow_queries"]


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


def _box(f, t0, centre, half, yaw, tilt):
    """12 triangles from t0 on: a box of half extents `half`, turned by `yaw` about y and `tilt` about x.  The eight
    corners are computed once, so neighbouring faces share them bit for bit (one object group); every face is a quad
    cut as (v0,v1,v2), (v0,v2,v3), the way scene.swift cuts them."""
    cy, sy, ct, st = np.cos(yaw), np.sin(yaw), np.cos(tilt), np.sin(tilt)
    R = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]]) @ np.array([[1, 0, 0], [0, ct, -st], [0, st, ct]])
    corner = {}
    for i in (-1, 1):
        for j in (-1, 1):
            for k in (-1, 1):
                corner[(i, j, k)] = (np.asarray(centre) + R @ (np.array([i, j, k]) * half)).astype(np.float32)
    faces = [[(-1, -1, -1), (-1, -1, 1), (-1, 1, 1), (-1, 1, -1)], [(1, -1, -1), (1, 1, -1), (1, 1, 1), (1, -1, 1)],
             [(-1, -1, -1), (1, -1, -1), (1, -1, 1), (-1, -1, 1)], [(-1, 1, -1), (-1, 1, 1), (1, 1, 1), (1, 1, -1)],
             [(-1, -1, -1), (-1, 1, -1), (1, 1, -1), (1, -1, -1)], [(-1, -1, 1), (1, -1, 1), (1, 1, 1), (-1, 1, 1)]]
    for q, (a, b, c, d) in enumerate(faces):
        for k, p in enumerate((a, b, c, a, c, d)):
            o = (3 * (t0 + 2 * q) + k) * 4
            f[o:o + 3] = corner[p]


def test_culling_never_changes_a_query_random_boxes():
    """Seeded fuzz (24 scenes): the two boxes of the Cornell scene replaced by boxes of random size, place, yaw and tilt (some poke
    through a wall, some cut each other, some are slivers), seen by a camera inside the room, four bounces: every
    query of every frame gives the full loop's answer."""
    W, H, spp = 120, 90, 4
    g = np.random.default_rng(11)
    total = 0
    for scene in range(24):
        sc = O.cornell(W, H)
        f = sc["vertices"].view(np.float32)
        for t0 in (10, 22):
            half = g.uniform(0.05, 1.1, 3) * (g.random(3) < 0.8) + 0.02
            _box(f, t0, g.uniform(-2.3, 2.3, 3), half, g.uniform(0, np.pi), g.uniform(-0.6, 0.6) * (scene % 2))
        cam = sc["camera"].view(np.float32)
        cam[0:3] = g.uniform(-1.5, 1.5, 3) + (0, 0, 1.0)
        cam[4:7] = g.uniform(-0.4, 0.4, 3) + (0, 0, -1.0)
        cam[14] = g.uniform(0.8, 1.6)
        grp = O.groups(sc)
        assert int(grp[2]) == 1 and 1 <= int(grp[0]) <= 3, (scene, grp[:4])
        r = O.cullcheck(sc, W, H, spp, bounces=4)
        assert r["mismatches"] == 0, (scene, r["mismatches"])
        total += r["queries"]
    assert total > 2_500_000


def test_culling_never_changes_a_query_modeb_frame():
    """Mode B (csrc/modeb.cuh: closest_b): the box rule per ray and the shell rule for the segments towards a light sample
    point, whose origins sit 1e-4 off their surface (mode A: 1e-3) -- every closest-hit query of a drawTriangle frame."""
    for W, H, ls in ((96, 64, 1.5), (64, 48, 1.0)):
        sc = O.cornell(W, H, light_size=ls)
        r = O.cullcheck_modeb(sc, W, H, camera_rays=2, mis_samples=60)
        _, ref, _ = O.drawtriangle(sc, W, H, camera_rays=2, mis_samples=60)
        assert np.array_equal(r["text"], ref)              # the hook does not disturb the oracle
        assert r["queries"] > 400_000 and r["mismatches"] == 0
        assert r["culled_tests"] < 0.3 * r["full_tests"]
        assert r["seg_in_shell"] > 0.999 * r["seg_queries"] > 0


def other_room(W, H, mode):
    """The Cornell scene in another room: "stretched" = every vertex and the light scaled and shifted by an affine map with
    three different factors (the room is no cube, nothing sits at a round number); "flush" = the light raised into the
    ceiling's plane (the ceiling is then one of the quads in the light's plane, and coplanar with the light's own)."""
    sc = O.cornell(W, H)
    f = sc["vertices"].view(np.float32).reshape(-1, 4)
    lt = sc["lights"].view(np.float32)
    if mode == "stretched":
        s = np.array([1.3, 0.7, 1.1], np.float32); t = np.array([0.4, 0.3, -0.6], np.float32)
        f[:, :3] = f[:, :3] * s + t
        lt[0:3] = lt[0:3] * s + t
    else:
        top = np.float32(2.5)
        f[f[:, 1] == np.float32(2.49), 1] = top
        lt[1] = top
    return sc


def test_exit_and_light_plane_rules_in_other_rooms():
    W, H, spp = 160, 120, 6
    for mode in ("stretched", "flush"):
        sc = other_room(W, H, mode)
        grp = O.groups(sc)
        flush = mode == "flush"
        # two boxes as groups; the shell: five walls -- and the light's quad when it lies in the ceiling's plane
        assert int(grp[2]) == 1 and int(grp[0]) == 2 and bin(int(grp[3])).count("1") == (6 if flush else 5)
        assert [bin(int(x)).count("1") for x in grp[44:50]] == [1, 1, 1, 2 if flush else 1, 1, 0]   # face by face; none in front
        assert bin(int(grp[50])).count("1") == (2 if flush else 1)                # in the light's plane: its quad; the ceiling too when flush
        for cam in (None, (0.3, -0.4, 1.5)):
            if cam is not None:
                c = sc["camera"].view(np.float32)
                c[0:3] = cam; c[4:7] = (0.1, -0.1, -1.0); c[14] = 1.4
            r = O.cullcheck(sc, W, H, spp, bounces=4)
            assert r["queries"] > 250_000 and r["mismatches"] == 0, (mode, cam, r["mismatches"])


def test_shell_rule_of_the_several_lights_passes_sphere_scene():
    """csrc/wave.cuh skips the shell triangles (the walls) for all shadow rays of a pass when every light's sample rectangle and
    every lane's shading point lie inside the shrunk shell box -- on config 4's scene too, whose triangles are grouped for
    the shell's sake only.  Every shadow query of a frame from inside the box, repeated without the walls but with all spheres."""
    W, H = 320, 180
    sc = O.spheres(W, H)
    r = O.shellcheck(sc, W, H, 4)
    g = r["groups"]
    assert int(g[2]) == 0 and int(g[3]) == 0b11111 and int(g[51]) == 1           # not for the sorted kernels; five walls; lights inside
    assert r["shadow"] > 1_500_000 and r["skipped"] > 0.999 * r["shadow"] and r["mismatches"] == 0
    assert np.array_equal(r["image"], O.pathtrace(sc, W, H, 4)[0])
    # a light that reaches out of the box switches the rule off (the word, not the geometry, decides)
    sc["lights"].view(np.float32)[0] = 2.4
    assert int(O.shellcheck(sc, 64, 36, 1)["groups"][51]) == 0
    # the Cornell scene: one light, inside
    assert int(O.groups(O.cornell(64, 64))[51]) == 1


def modeb_scenes(W, H):
    """Scenes for mode B's group / shell rule away from the Cornell box: the two rooms of other_room, and four seeded scenes
    whose boxes have random size, place, yaw and tilt (some through a wall), each with a camera inside the room."""
    out = []
    for mode in ("stretched", "flush"):
        out.append((mode, other_room(W, H, mode)))
    g = np.random.default_rng(23)
    for k in range(4):
        sc = O.cornell(W, H, light_size=1.5)
        f = sc["vertices"].view(np.float32)
        for t0 in (10, 22):
            _box(f, t0, g.uniform(-2.3, 2.3, 3), g.uniform(0.05, 1.1, 3) + 0.02, g.uniform(0, np.pi), g.uniform(-0.6, 0.6) * (k % 2))
        cam = sc["camera"].view(np.float32)
        cam[0:3] = g.uniform(-1.5, 1.5, 3) + (0, 0, 1.0); cam[4:7] = g.uniform(-0.4, 0.4, 3) + (0, 0, -1.0); cam[14] = g.uniform(0.8, 1.6)
        out.append((f"boxes{k}", sc))
    return out


def test_culling_never_changes_a_query_modeb_other_scenes():
    W, H = 64, 48
    total = 0
    for name, sc in modeb_scenes(W, H):
        r = O.cullcheck_modeb(sc, W, H, camera_rays=2, mis_samples=30)
        assert r["mismatches"] == 0, (name, r["mismatches"])
        total += r["queries"]
    assert total > 800_000
