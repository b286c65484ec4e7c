"""RT_VARIANT_MIS: the mode-A estimator with multiple importance sampling, on the CPU checker.

The reference's pathTrace weights nothing (its TODO: the commented-out block of RTrace/sampling.metal:238-260).  With the
flag, the direct light of a shaded bounce is estimated by the light sample and by the cosine-sampled continuation ray,
combined with the balance heuristic (oracle/rt_oracle.c states the estimator).  What pins it here:
  * the two strategies ALONE (checker-only selectors) and their MIS combination have the same expectation -- a wrong
    normalisation of the cosine term (pi / A_s), a wrong rectangle test or weights that do not add up to one break this;
  * the weights are powerHeuristic / calculateSquareLightPdf / calculateCosineWeightedPdf of shaders.metal:132-137,
    315-326, 376-380, checked against a hand computation in numpy fp32;
  * flag off, nothing moves (the reference-pinned tests run without it);
  * which kernel form the library picks with the flag (csrc/rt_plan.cpp)."""
import numpy as np
import pytest

import oracle_lib as O
import gpuraytracer_b200 as G
from gpuraytracer_b200 import _lib as L
from gpuraytracer_b200.api import choose_kernel, default_params

F = np.float32


def test_light_only_cosine_only_and_mis_have_one_expectation():
    W, H = 64, 48
    sc = O.cornell(W, H)
    light = O.pathtrace(sc, W, H, 1024, mis=O.MIS_LIGHT_ONLY)[0][..., :3].astype(np.float64)
    mis = O.pathtrace(sc, W, H, 1024, mis=O.MIS)[0][..., :3].astype(np.float64)
    cos = O.pathtrace(sc, W, H, 4096, mis=O.MIS_COSINE_ONLY)[0][..., :3].astype(np.float64)     # 500 x the variance: more samples
    off = O.pathtrace(sc, W, H, 1024)[0][..., :3].astype(np.float64)
    blocks = lambda a: a.reshape(6, 8, 8, 8, 3).mean(axis=(1, 3))
    # MIS against the light strategy alone: per 8 x 8 block (both are low-noise estimators of the same image)
    rel = np.abs(blocks(mis) - blocks(light)) / np.maximum(blocks(light), 1e-3)
    assert rel.max() < 0.03 and rel.mean() < 0.004, (rel.max(), rel.mean())
    # the cosine strategy alone is noisy (it must hit a 0.5 x 0.5 rectangle): global means per channel
    g = lambda a: a.mean(axis=(0, 1))
    assert np.all(np.abs(g(cos) - g(light)) / g(light) < 0.03), (g(cos), g(light))
    assert np.all(np.abs(g(mis) - g(light)) / g(light) < 0.002), (g(mis), g(light))
    # and the flag-off image is ANOTHER image: raytrace.metal:59's `accumulated = emissive` after a bounce is not an
    # estimate of the same integral (documented in rt_oracle.c), so the test must be able to tell images apart
    assert np.all(np.abs(g(off) - g(light)) / g(light) > 0.2)


def test_weights_by_hand():
    """The light strategy's weight, read off the image (bounces = 1: acc = w_l * g, plus the cosine term on the few paths
    whose continuation ray goes through the light's rectangle), then the closed formula on one point in fp32."""
    W, H = 32, 24
    sc = O.cornell(W, H)
    a = O.pathtrace(sc, W, H, 1, bounces=1, mis=O.MIS_LIGHT_ONLY)[0][..., 0]
    b = O.pathtrace(sc, W, H, 1, bounces=1, mis=O.MIS)[0][..., 0]
    c = O.pathtrace(sc, W, H, 1, bounces=1, mis=O.MIS_COSINE_ONLY)[0][..., 0]
    lit = (a > 1e-6) & (a < 0.99) & (c == 0)                           # shaded, lit, no cosine-strategy term, not the light itself
    assert lit.sum() > 100
    ratio = b[lit] / a[lit]                                             # = w_l
    assert ratio.max() <= 1.0 + 1e-6 and ratio.min() > 0.5 and np.median(ratio) > 0.9, (ratio.min(), np.median(ratio), ratio.max())
    # exact form on a synthetic point, in fp32, against the closed formula
    P, n = np.array([0.3, -2.499, 0.4], F), np.array([0, 1, 0], F)
    c, nL = np.array([0, 2.49, 0], F), np.array([0, -1, 0], F)
    d = (c - P) / F(np.linalg.norm((c - P).astype(np.float64)))
    p_l = F(np.dot(c - P, c - P)) / (F(1.0) * max(F(0), F(np.dot(-d, nL))) + F(1e-6))
    p_c = max(F(0), F(np.dot(n, d))) / F(np.pi)
    w_l, w_c = p_l / (p_l + p_c + F(1e-6)), p_c / (p_c + p_l + F(1e-6))
    assert abs(float(w_l + w_c) - 1.0) < 1e-6 and 0.98 < w_l < 1.0


def test_light_normals_table_default_is_the_reference():
    """A table holding (0,-1,0) gives the reference's image up to the sign of a zero (the general path multiplies by the basis
    instead of folding the zeros); a tilted normal gives another image."""
    W, H = 48, 36
    sc = O.cornell(W, H)
    ref = O.pathtrace(sc, W, H, 8)[0]
    same = O.pathtrace(sc, W, H, 8, light_normals=np.array([[0, -1, 0, 0]], F))[0]
    assert np.allclose(same, ref, rtol=1e-6, atol=1e-7)
    tilted = O.pathtrace(sc, W, H, 8, light_normals=np.array([[0.6, -0.8, 0, 0]], F))[0]
    assert not np.allclose(tilted, ref, rtol=1e-2)
    assert np.isfinite(tilted).all()


def test_plan_with_mis_flag():
    cornell = lambda **kw: default_params(1920, 1080, row0=0, rows=1080, numTriangles=36, numLights=1, **kw)
    name = lambda p: choose_kernel(p)["kernel"]
    assert name(cornell(spp=1024, variant=L.VARIANT_MIS)) == "pathTrace_sortedblk6_mis"
    assert name(cornell(spp=1024, variant=L.VARIANT_MIS | L.VARIANT_SMEM)) == "pathTrace_smem_mis"
    assert name(cornell(spp=1024, variant=L.VARIANT_MIS | L.VARIANT_CONST)) == "pathTrace_smem_mis"
    assert name(cornell(spp=1024, variant=L.VARIANT_MIS | L.VARIANT_WAVE)) == "pathTrace_wave_mis"
    assert name(cornell(spp=1024, variant=L.VARIANT_MIS | L.VARIANT_TILES)) == "pathTrace_wave_mis"
    assert name(cornell(spp=1024, variant=L.VARIANT_MIS | L.VARIANT_SORTED)) == "pathTrace_wave_mis"
    assert name(cornell(spp=1024, variant=L.VARIANT_MIS | L.VARIANT_SORTED_BLOCK_TILES)) == "pathTrace_sortedblk6_mis"
    assert name(cornell(spp=1024, variant=L.VARIANT_MIS | L.VARIANT_BVH)) == "pathTrace_bvh_mis"
    assert name(cornell(spp=2, variant=L.VARIANT_MIS)) == "pathTrace_wave_mis"                 # progressive: any sample count
    sph = default_params(1920, 1080, row0=0, rows=1080, numTriangles=18, numSpheres=256, numLights=4, spp=256, variant=L.VARIANT_MIS)
    assert name(sph) == "pathTrace_wave4_mis"
    with pytest.raises(G.RTError):
        name(cornell(spp=64, variant=L.VARIANT_MIS | L.VARIANT_COUNT_RAYS))
