"""GPU parity of RT_VARIANT_MIS and of the per-light normal table against the CPU checker, bit for bit,
through the C ABI: every kernel form that carries the flag (per-pixel, compacting R = 6 and R = 4, block-sorted, BVH),
the forms that fall back to them, several lights, and lights whose normals are not the reference's (0,-1,0)."""
import numpy as np
import pytest

import gpuraytracer_b200 as G
from gpuraytracer_b200 import _lib as L
import oracle_lib as O

pytestmark = pytest.mark.gpu
F = np.float32


@pytest.fixture(scope="module")
def dev():
    return G.Device(0)


@pytest.mark.parametrize("bounces", [1, 3])
def test_mis_every_form_cornell(dev, bounces):
    W, H, spp = 200, 77, 13                                     # ragged against the block; odd sample count
    ref, _ = O.pathtrace(O.cornell(W, H), W, H, spp, bounces=bounces, mis=O.MIS)
    off, _ = O.pathtrace(O.cornell(W, H), W, H, spp, bounces=bounces)
    assert not np.array_equal(ref, off)                         # the flag changes the image
    r = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev)
    for name in ("DEFAULT", "SMEM", "CONST", "WAVE", "TILES", "WIDE", "SORTED", "SORTED_TILES", "SORTED_BLOCK", "SORTED_BLOCK_TILES", "BVH"):
        img = r.draw(spp=spp, bounces=bounces, variant=getattr(L, "VARIANT_" + name) | L.VARIANT_MIS)
        assert np.array_equal(img, ref), (name, int((img != ref).sum()))
    # flag off again on the same renderer: the reference's image, untouched
    assert np.array_equal(r.draw(spp=spp, bounces=bounces), off)


def test_mis_progressive_and_few_samples(dev):
    W, H = 96, 64
    ref, _ = O.pathtrace(O.cornell(W, H), W, H, 10, mis=O.MIS)
    r = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev)
    r.clear()
    r.draw_async(spp=3, firstSample=0, accumulate=1, variant=L.VARIANT_MIS)
    r.draw_async(spp=7, firstSample=3, accumulate=1, variant=L.VARIANT_MIS).wait_until_completed()
    acc = r.image()
    assert np.all(acc[..., 3] == 10) and np.array_equal(acc[..., :3] / F(10), ref[..., :3])


def test_mis_spheres_four_lights_and_extent(dev):
    W, H, spp = 96, 64, 5
    osc = O.spheres(W, H)
    r = G.Renderer(G.init_sphere_scene(W, H), W, H, device=dev)
    for extent in (False, True):
        ref, _ = O.pathtrace(osc, W, H, spp, mis=O.MIS, light_extent=extent)
        flags = L.VARIANT_MIS | (L.VARIANT_LIGHT_EXTENT if extent else 0)
        for name in ("DEFAULT", "SMEM", "WAVE"):
            img = r.draw(spp=spp, variant=getattr(L, "VARIANT_" + name) | flags)
            assert np.array_equal(img, ref), (extent, name, int((img != ref).sum()))


def test_mis_bvh_tessellated(dev):
    from bvh_util import tessellate
    W, H, spp = 80, 60, 4
    sc = G.init_cornell_box(W, H)
    osc = O.cornell(W, H)
    tv, tm, T = tessellate(sc.vertices, sc.materials, 36, 6)
    sc.vertices, sc.materials, sc.num_triangles = tv, tm, T
    osc = dict(osc, vertices=tv, materials=tm, T=T)
    ref, _ = O.pathtrace(osc, W, H, spp, mis=O.MIS)
    img = G.Renderer(sc, W, H, device=dev).draw(spp=spp, variant=L.VARIANT_MIS)
    assert np.array_equal(img, ref), int((img != ref).sum())


@pytest.mark.parametrize("mis", [0, 1])
def test_light_normal_table(dev, mis):
    W, H, spp = 120, 90, 6
    flag_o, flag_g = (O.MIS, L.VARIANT_MIS) if mis else (0, 0)
    # one light, tilted; then the reference's own normal given through the table (the general path: same image up to the
    # sign of zeros -- compared with the checker run the same way, bit for bit)
    for normal in ([0.6, -0.8, 0.0, 0.0], [0.0, -1.0, 0.0, 0.0], [-0.3, -0.5, 0.4, 0.0]):
        n = np.array([normal], F)
        ref, _ = O.pathtrace(O.cornell(W, H), W, H, spp, mis=flag_o, light_normals=n)
        r = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev)
        r.set_light_normals(n)
        for name in ("DEFAULT", "SMEM", "WAVE", "SORTED_BLOCK", "BVH"):
            img = r.draw(spp=spp, variant=getattr(L, "VARIANT_" + name) | flag_g)
            assert np.array_equal(img, ref), (normal, name, int((img != ref).sum()))
        # unbinding the table brings the reference's literals back
        r.set_light_normals(None)
        assert np.array_equal(r.draw(spp=spp, variant=flag_g), O.pathtrace(O.cornell(W, H), W, H, spp, mis=flag_o)[0])
    # four lights, each its own normal, with the light-extent flag
    osc = O.spheres(W, H)
    n4 = np.array([[0, -1, 0, 0], [0.5, -0.8, 0.1, 0], [-0.4, -0.9, 0, 0], [0, -0.7, -0.7, 0]], F)
    ref, _ = O.pathtrace(osc, W, H, 4, mis=flag_o, light_normals=n4, light_extent=True)
    r = G.Renderer(G.init_sphere_scene(W, H), W, H, device=dev)
    r.set_light_normals(n4)
    for name in ("DEFAULT", "SMEM"):
        img = r.draw(spp=4, variant=getattr(L, "VARIANT_" + name) | flag_g | L.VARIANT_LIGHT_EXTENT)
        assert np.array_equal(img, ref), (name, int((img != ref).sum()))
