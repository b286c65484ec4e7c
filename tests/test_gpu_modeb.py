"""GPU parity of the MIS kernel (mode B, `drawTriangle`, Sources/gpuRaytracer/shaders.metal:635-707)
through the C ABI against the oracle and against the golden vectors made from the reference's own
shader source.  Parity is judged on the float3 buffer (bit equality demanded, 1e-4 relative is the
north_star bar); the RGBA8 image may differ by at most 1 LSB."""
import os

import numpy as np
import pytest

import oracle_lib as O
import gpuraytracer_b200 as G

pytestmark = pytest.mark.gpu
REL_TOL = 1e-4
G_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def check(text, pixels, want_text, want_pixels):
    bad = int((np.abs(text - want_text) > REL_TOL * np.maximum(np.abs(want_text), 1e-6)).sum())
    assert bad == 0, f"{bad} values out of 1e-4 relative"
    assert np.array_equal(text.view(np.uint32), want_text.view(np.uint32)), f"{(text != want_text).sum()} values differ"
    assert int(np.abs(pixels.astype(int) - want_pixels.astype(int)).max()) <= 1


def test_modeb_matches_golden_from_reference_shader():
    g = np.load(os.path.join(G_DIR, "drawtriangle_cornell_64x48_ref.npz"))
    px, tx = G.draw_triangle(G.init_cornell_box(64, 48, light_size=1.5), 64, 48)     # the shader's literals: 6 x 300
    check(tx, px, g["text"], g["pixels"])


def test_modeb_config1_512(dev=None):
    """BASELINE.json configs[0] size for the MIS kernel: Cornell 512x512 (oracle on a row band, GPU on the frame)."""
    W = H = 512
    px, tx, cmd = G.draw_triangle(G.init_cornell_box(W, H, light_size=1.5), W, H, return_command=True)
    assert cmd.launch_count() == 2 and cmd.gpu_time_ms() > 0
    want_px, want_tx, _ = O.drawtriangle(O.cornell(W, H, 1.5), W, H, row0=240, rows=24)
    check(tx[240:264], px[240:264], want_tx, want_px)
    assert not np.isnan(tx).any() and tx.max() > 1.0


def test_modeb_parameters_and_row_tile():
    W, H = 96, 72
    sc = G.init_cornell_box(W, H, light_size=1.5)
    px, tx = G.draw_triangle(sc, W, H, row0=16, rows=8, camera_rays=2, mis_samples=30)
    want_px, want_tx, _ = O.drawtriangle(O.cornell(W, H, 1.5), W, H, row0=16, rows=8, camera_rays=2, mis_samples=30)
    check(tx, px, want_tx, want_px)
    # interleaved tiles (multi-GPU deal): local rows 0..7 = global rows 4..7 and 12..15
    px2, tx2 = G.draw_triangle(sc, W, H, row0=4, rows=8, camera_rays=2, mis_samples=30, tile_rows=4, tile_stride=8)
    w2p, w2t, _ = O.drawtriangle(O.cornell(W, H, 1.5), W, H, row0=4, rows=8, camera_rays=2, mis_samples=30,
                                 tile_rows=4, tile_stride=8)
    check(tx2, px2, w2t, w2p)


def test_modeb_group_and_shell_rule_other_scenes():
    """closest_b's group rule over a warp's union and its shell rule (origins 1e-4 off their surface) away from the Cornell
    box: tests/test_group_cull_cpu.py: modeb_scenes -- the CPU checker's side of the same scenes."""
    from test_group_cull_cpu import modeb_scenes
    W, H = 64, 48
    for name, osc in modeb_scenes(W, H):
        gsc = G.init_cornell_box(W, H, light_size=1.5)
        gsc.vertices[:] = osc["vertices"].view(np.uint8); gsc.lights[:] = osc["lights"].view(np.uint8); gsc.camera[:] = osc["camera"].view(np.uint8)
        px, tx = G.draw_triangle(gsc, W, H, camera_rays=2, mis_samples=30)
        want_px, want_tx, _ = O.drawtriangle(osc, W, H, camera_rays=2, mis_samples=30)
        assert np.array_equal(tx.view(np.uint32), want_tx.view(np.uint32)), (name, int((tx != want_tx).sum()))
