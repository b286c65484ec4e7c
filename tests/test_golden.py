"""Golden vectors generated from the reference's own shader source (tests/golden/make_golden.py).
CPU: the oracle reproduces them bit for bit.  GPU: so does the CUDA path through the C ABI --
that is the kernel checked against the reference's source directly, not through our restatement."""
import os

import numpy as np
import pytest

import oracle_lib as O

G_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden(name):
    return np.load(os.path.join(G_DIR, name))


def test_oracle_halton_matches_golden():
    g = golden("halton_ref.npz")
    got = np.array([O.lib.rt_oracle_halton(int(i), int(d)) for i, d in zip(g["i"], g["d"])], np.float32)
    assert np.array_equal(got.view(np.uint32), g["value"].view(np.uint32))


def test_oracle_pathtrace_matches_golden_image():
    g = golden("pathtrace_cornell_64x48_ref400spp.npz")
    sc = O.cornell(64, 48)
    for k in ("camera", "materials", "lights", "vertices"):
        assert np.array_equal(sc[k], g[k]), k            # the fixture bytes have not drifted
    got, _ = O.pathtrace(sc, 64, 48, 400, bounces=3, off=g["offsets"].copy())
    assert np.array_equal(got.view(np.uint32), g["image"].view(np.uint32))


def test_oracle_pathtrace_matches_golden_image_moved_camera():
    """The reference kernel seen from inside the room (off-axis, rolled, 100 degrees wide): 80 x 60 at its literals."""
    g = golden("pathtrace_cornell_80x60_moved_ref400spp.npz")
    sc = O.cornell(80, 60)
    for k in ("materials", "lights", "vertices"):
        assert np.array_equal(sc[k], g[k]), k
    sc["camera"] = g["camera"].copy()
    got, _ = O.pathtrace(sc, 80, 60, 400, bounces=3, off=g["offsets"].copy())
    assert np.array_equal(got.view(np.uint32), g["image"].view(np.uint32))


def test_oracle_drawtriangle_matches_golden():
    g = golden("drawtriangle_cornell_64x48_ref.npz")
    sc = O.cornell(64, 48, light_size=1.5)
    for k in ("camera", "materials", "lights", "vertices"):
        assert np.array_equal(sc[k], g[k]), k
    px, tx, _ = O.drawtriangle(sc, 64, 48)
    assert np.array_equal(tx.view(np.uint32), g["text"].view(np.uint32))
    assert np.array_equal(px, g["pixels"])


@pytest.mark.gpu
def test_gpu_pathtrace_matches_golden_image():
    import gpuraytracer_b200 as G
    g = golden("pathtrace_cornell_64x48_ref400spp.npz")
    r = G.Renderer(G.init_cornell_box(64, 48), 64, 48, offsets=g["offsets"].copy())
    img = r.draw()                                        # no overrides: the shader's own 400 spp / 3 bounces
    assert np.array_equal(img.view(np.uint32), g["image"].view(np.uint32)), f"{(img != g['image']).sum()} values differ"


@pytest.mark.gpu
def test_gpu_pathtrace_matches_golden_image_moved_camera():
    import gpuraytracer_b200 as G
    g = golden("pathtrace_cornell_80x60_moved_ref400spp.npz")
    scene = G.init_cornell_box(80, 60)
    scene.camera.view(np.uint8)[:] = g["camera"].view(np.uint8)
    r = G.Renderer(scene, 80, 60, offsets=g["offsets"].copy())
    img = r.draw()
    assert np.array_equal(img.view(np.uint32), g["image"].view(np.uint32)), f"{(img != g['image']).sum()} values differ"


@pytest.mark.gpu
def test_gpu_halton_matches_golden():
    import ctypes as C
    import gpuraytracer_b200 as G
    from gpuraytracer_b200 import _lib as L
    g = golden("halton_ref.npz")
    pairs = np.stack([g["i"], g["d"]], 1).astype(np.uint32).copy()
    dev = G.Device(0)
    for what in (0, 3):
        out = np.zeros(pairs.shape[0], np.float32)
        L.check(L.lib.rtDebugEvaluate(dev.h, what, pairs.ctypes.data, pairs.nbytes, out.ctypes.data, out.nbytes, out.shape[0]),
                "rtDebugEvaluate")
        assert np.array_equal(out.view(np.uint32), g["value"].view(np.uint32)), what
