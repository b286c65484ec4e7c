"""The sphere grid (csrc/sgrid.cuh) on the device: the kernels that walk it against the per-pixel kernel, which never
does, at config 4's full frame size, and against the oracle on the rows that once differed.  The CPU side of the same
argument is test_sgrid_cpu.py."""
import numpy as np
import pytest

import gpuraytracer_b200 as G
from gpuraytracer_b200 import _lib as L
import oracle_lib as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    return G.Device(0)


def test_config4_full_frame_every_form(dev):
    """1920 x 1080 x 4 spp, 256 spheres + 4 lights: every form that walks the grid gives the per-pixel kernel's frame (the
    first form of the walk lost ONE pixel of these two million), and rows 672..687 are the oracle's."""
    W, H, spp = 1920, 1080, 4
    r = G.Renderer(G.init_sphere_scene(W, H), W, H, device=dev)
    base = r.draw(spp=spp, variant=L.VARIANT_SMEM)
    ref, _ = O.pathtrace(O.spheres(W, H), W, H, spp, row0=672, rows=16)
    assert np.array_equal(base[672:688], ref)
    for variant in (L.VARIANT_DEFAULT, L.VARIANT_WAVE, L.VARIANT_TILES, L.VARIANT_WIDE):
        img = r.draw(spp=spp, variant=variant)
        bad = np.argwhere((img != base).any(axis=2))
        assert len(bad) == 0, (variant, len(bad), bad[:4].tolist())


def test_other_sphere_counts_and_a_crowded_cell(dev):
    """16 and 64 spheres; then twelve spheres on one spot: the cell overflows, the build flags the grid unusable and the
    kernels fall back to the loop over all spheres -- same image either way."""
    W, H = 160, 90
    for n in (16, 64):
        ref, _ = O.pathtrace(O.spheres(W, H, n=n), W, H, 4)
        r = G.Renderer(G.init_sphere_scene(W, H, num_spheres=n), W, H, device=dev)
        for variant in (L.VARIANT_SMEM, L.VARIANT_WAVE, L.VARIANT_DEFAULT):
            assert np.array_equal(r.draw(spp=4, variant=variant), ref), (n, variant)
    osc = O.spheres(W, H, n=64); gsc = G.init_sphere_scene(W, H, num_spheres=64)
    osc["spheres"].view(np.float32).reshape(-1, 20)[:12, 0:3] = osc["spheres"].view(np.float32).reshape(-1, 20)[0, 0:3]
    gsc.spheres.view(np.float32).reshape(-1, 20)[:12, 0:3] = gsc.spheres.view(np.float32).reshape(-1, 20)[0, 0:3]
    assert not O.sgridcheck(osc, W, H, 1)["ok"]
    ref, _ = O.pathtrace(osc, W, H, 4)
    r = G.Renderer(gsc, W, H, device=dev)
    for variant in (L.VARIANT_SMEM, L.VARIANT_WAVE, L.VARIANT_DEFAULT):
        assert np.array_equal(r.draw(spp=4, variant=variant), ref), variant


def test_random_sphere_scenes(dev):
    """Spheres thrown at random into config 4's room (tests/test_sgrid_cpu.py: random_sphere_scene -- radii from 0.02 to 0.9,
    some through a wall, every tenth on its neighbour's centre; the crowded scenes overflow a cell and fall back to the full
    loop on the device as they do in the checker): every form against the oracle, four bounces."""
    from test_sgrid_cpu import random_sphere_scene
    W, H = 120, 68
    for seed, n in ((1, 1), (2, 7), (3, 40), (4, 40), (5, 120), (6, 256), (7, 256), (8, 19)):
        osc = random_sphere_scene(W, H, seed, n)
        gsc = G.init_sphere_scene(W, H, num_spheres=n)
        gsc.spheres[:] = osc["spheres"].view(np.uint8)
        ref, _ = O.pathtrace(osc, W, H, 4, bounces=4)
        r = G.Renderer(gsc, W, H, device=dev)
        for variant in (L.VARIANT_SMEM, L.VARIANT_DEFAULT, L.VARIANT_WAVE, L.VARIANT_TILES):
            img = r.draw(spp=4, bounces=4, variant=variant)
            assert np.array_equal(img, ref), (seed, n, variant, int((img != ref).any(axis=2).sum()))


def test_lights_off_the_common_plane_or_out_of_the_shell_box(dev):
    """The several-lights passes drop the walls and the lights' own quads from their shadow rays only when rt_prep_scene says
    they may (RT_GRP_LIGHTS_IN, RT_GRP_LIGHT_PLANE).  A light lowered out of the others' plane, a light whose sample
    rectangle reaches through the wall, both: the rules switch off piecewise and the image stays the oracle's."""
    W, H = 160, 90
    for case in ("lowered", "through_wall", "both"):
        osc = O.spheres(W, H, n=64); gsc = G.init_sphere_scene(W, H, num_spheres=64)
        for lights in (osc["lights"].view(np.float32), gsc.lights.view(np.float32)):
            if case in ("lowered", "both"):
                lights[1] = 2.0                     # light 0: centre.y
            if case in ("through_wall", "both"):
                lights[16 + 0] = 2.4                # light 1: centre.x, 0.1 from the wall, half extent 0.25
        g = O.shellcheck(osc, 32, 18, 1)["groups"]
        assert int(g[50]) == (0 if case != "through_wall" else 0b111100000) and int(g[51]) == (1 if case == "lowered" else 0)
        ref, _ = O.pathtrace(osc, W, H, 4)
        r = G.Renderer(gsc, W, H, device=dev)
        for variant in (L.VARIANT_SMEM, L.VARIANT_DEFAULT, L.VARIANT_WAVE):
            assert np.array_equal(r.draw(spp=4, variant=variant), ref), (case, variant)
