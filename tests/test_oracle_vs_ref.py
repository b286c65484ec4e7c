"""Pins the C restatement (oracle/rt_oracle.c) against the reference's own shader source:
oracle/_ref is RTrace/raytrace.metal + sampling.metal, unmodified, compiled as C++.  The two
must agree BIT FOR BIT at the shader's literals (400 spp, 3 bounces).  What this cannot pin
is the closed-source part both share by definition (oracle/rt_detmath.h): Metal's library
functions and its ray/triangle query."""
import numpy as np
import pytest

import oracle_lib as O
import ref_lib as R

ref = R.load()
pytestmark = pytest.mark.skipif(ref is None, reason="oracle/_ref not built (needs /root/reference)")


def test_halton_and_hash_match_reference_source():
    rng = np.random.default_rng(7)
    for i in list(range(0, 64)) + [int(v) for v in rng.integers(0, 1 << 21, 400)] + [0xffffffff, 1 << 24]:
        for d in range(24):
            a, b = O.lib.rt_oracle_halton(i, d), ref.rt_ref_halton(i, d)
            assert np.float32(a).view(np.uint32) == np.float32(b).view(np.uint32), (i, d)
    for x in [0, 1, 2, 0xdeadbeef, 0xffffffff] + [int(v) for v in rng.integers(0, 1 << 32, 200)]:
        assert O.lib.rt_oracle_hash(x) == ref.rt_ref_hash(x)


@pytest.mark.parametrize("wh", [(64, 48), (96, 96)])
def test_pathtrace_restatement_is_bit_identical_to_reference_shader(wh):
    W, H = wh
    sc = O.cornell(W, H)
    off = O.offsets(W, 0, H, seed=3)
    want = R.pathtrace(ref, sc, W, H, off)
    got, cnt = O.pathtrace(sc, W, H, 400, bounces=3, off=off)
    assert int(cnt[0]) == W * H * 400
    assert want[..., 3].min() == 1.0 and want[..., :3].max() > 0.5          # an image, not zeros
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), f"{(got != want).sum()} values differ"


def test_reference_row_window_matches_full_frame():
    W, H = 64, 48
    sc = O.cornell(W, H)
    off = O.offsets(W, 0, H, seed=5)
    part = R.pathtrace(ref, sc, W, H, off, row0=20, rows=4)
    got, _ = O.pathtrace(sc, W, H, 400, row0=20, rows=4, off=off[20 * W:24 * W].copy())
    assert np.array_equal(got, part)


# ---- mode B: drawTriangle (MIS), Sources/gpuRaytracer/shaders.metal ---------------------------
def test_drawtriangle_restatement_is_bit_identical_to_reference_shader():
    W, H = 64, 48
    sc = O.cornell(W, H, light_size=1.5)          # Sources/gpuRaytracer/main.swift:30-31
    want_px, want_tx = R.drawtriangle(ref, sc, W, H)
    got_px, got_tx, cnt = O.drawtriangle(sc, W, H)   # defaults = the shader's literals: 6 rays, 300 MIS samples
    assert int(cnt[0]) == W * H * 6
    assert want_tx.max() > 1.0 and not np.isnan(want_tx).any()
    assert np.array_equal(got_tx.view(np.uint32), want_tx.view(np.uint32)), f"{(got_tx != want_tx).sum()} values differ"
    assert np.array_equal(got_px, want_px)


def test_drawtriangle_4x3_aspect_and_hash_random():
    """800x600-like aspect (the shader hard-codes 800/600 in hashRandom): a small 4:3 window, a few rows."""
    import ctypes as C
    W, H = 80, 60
    sc = O.cornell(W, H, light_size=1.5)
    want_px, want_tx = R.drawtriangle(ref, sc, W, H, row0=28, rows=3)
    got_px, got_tx, _ = O.drawtriangle(sc, W, H, row0=28, rows=3)
    assert np.array_equal(got_tx.view(np.uint32), want_tx.view(np.uint32))
    assert np.array_equal(got_px, want_px)
    out = (C.c_float * 2)()
    ref.rt_ref_b_hash_random(5, 7, 3, out)
    assert 0.0 <= out[0] <= 1.0 and 0.0 <= out[1] <= 1.0
