"""No-GPU checks of the product boundary: the C-ABI library loads, exports every
symbol include/rtcuda.h declares, reports errors as codes, and the C++ host mirror
produces the same scene bytes as the oracle's independent restatement."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import oracle_lib as O
import gpuraytracer_b200 as G
from gpuraytracer_b200 import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_every_declared_symbol_is_exported():
    hdr = open(os.path.join(ROOT, "include", "rtcuda.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(rt[A-Z]\w+)\s*\(", hdr))
    assert len(declared) >= 40
    raw = C.CDLL(L.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(raw, name), f"librtcuda.so lacks {name}"
    assert declared == set(L.SIGNATURES), declared ^ set(L.SIGNATURES)
    assert L.lib.rtAbiVersion() == 1


def test_python_constants_mirror_the_header_enums():
    """Every RT_INDEX_*, RT_FORMAT_* and RT_VARIANT_* enumerator of include/rtcuda.h has the same value in _lib.py."""
    hdr = open(os.path.join(ROOT, "include", "rtcuda.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    enums = dict(re.findall(r"\bRT_((?:INDEX|FORMAT|VARIANT)_[A-Z0-9_]+)\s*=\s*(0x[0-9a-fA-F]+|\d+)", hdr))
    assert len(enums) >= 25
    for name, value in enums.items():
        if name == "INDEX_COUNT":
            continue
        assert getattr(L, name) == int(value, 0), name
    mine = {n for n in dir(L) if n.startswith(("INDEX_", "FORMAT_", "VARIANT_"))}
    assert mine <= set(enums), mine - set(enums)


def test_struct_sizes_match_header():
    assert C.sizeof(L.RTLaunchParams) == 64 and C.sizeof(O.RTLaunchParams) == 64
    assert C.sizeof(L.RTDeviceInfo) == 128 + 6 * 4 + 8


def test_default_params_are_the_reference_literals():
    p = G.default_params(800, 600)
    assert (p.width, p.height, p.row0, p.rows) == (800, 600, 0, 600)
    assert p.spp == 400 and p.bounces == 3          # raytrace.metal:24-25
    assert p.cameraRays == 6 and p.misSamples == 300  # shaders.metal:644,648
    assert L.lib.rtDefaultParams(0, 10, C.byref(p)) == -1


def test_argument_errors_are_codes_not_crashes():
    h = C.c_void_p()
    assert L.lib.rtDeviceCreate(-1, C.byref(h)) == -1
    assert L.lib.rtDeviceCreate(0, None) == -1
    assert L.lib.rtPipelineCreate(None, b"pathTrace", C.byref(h)) == -1
    assert L.lib.rtCommit(None) == -1 and L.lib.rtWaitUntilCompleted(None) == -1
    assert L.lib.rtBufferContents(None) is None
    assert b"NULL" in L.lib.rtLastErrorString()
    assert L.lib.rtBufferRelease(None) == 0 and L.lib.rtDeviceRelease(None) == 0


@pytest.mark.skipif(os.path.exists("/dev/nvidiactl"), reason="a GPU driver is present")
def test_no_driver_fails_loudly_without_fallback():
    with pytest.raises(G.RTError) as e:
        G.Device(0)
    assert e.value.code == -3 and "libcuda" in str(e.value)


def test_host_mirror_bytes_equal_oracle_fixture():
    for w, h, ls in ((800, 600, 1.0), (1920, 1080, 1.0), (512, 512, 1.5)):
        s = G.init_cornell_box(w, h, ls)
        f = O.cornell(w, h, ls)
        assert (s.num_triangles, s.num_lights, s.num_spheres) == (36, 1, 0)
        for k in ("camera", "materials", "lights", "vertices"):
            assert np.array_equal(getattr(s, k), f[k]), k
    s = G.init_sphere_scene(320, 200, 12345, 256)
    f = O.spheres(320, 200, 12345, 256)
    assert (s.num_triangles, s.num_lights, s.num_spheres) == (18, 4, 256)
    for k in ("camera", "materials", "lights", "vertices", "spheres"):
        assert np.array_equal(getattr(s, k), f[k]), k
    assert np.array_equal(G.fill_offsets(200, 7, 33, 5), O.offsets(200, 7, 33, 5))
    assert G.fill_offsets(64, 0, 64).max() < (1 << 20)


def test_sphere_scene_is_sane():
    s = G.init_sphere_scene(64, 64, 12345, 256)
    sp = s.spheres.view(np.float32).reshape(256, 20)
    c, r = sp[:, :3], sp[:, 16]
    assert r.min() >= 0.1 and r.max() < 0.2
    assert np.all(np.abs(c) + r[:, None] < 2.5)          # inside the room
    d = np.linalg.norm(c[:, None] - c[None], axis=-1) + np.eye(256) * 9
    assert np.all(d > r[:, None] + r[None])              # no two spheres overlap


def test_png_writer_of_the_cpp_host_round_trips(tmp_path):
    """image.swift:102-157 mirror (csrc/host_renderer.cpp): RGBA8 -> PNG with stored deflate blocks, read back by PIL."""
    import ctypes as C
    from PIL import Image
    from gpuraytracer_b200._lib import lib
    rng = np.random.default_rng(5)
    for (w, h) in ((1, 1), (37, 19), (800, 100)):       # 800x100x4+100 bytes: more than one 65535-byte deflate block
        px = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
        path = str(tmp_path / f"t_{w}x{h}.png").encode()
        lib.rtHostSavePixelArrayToImage.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_char_p]
        assert lib.rtHostSavePixelArrayToImage(px.ctypes.data, w, h, path) == 0
        back = np.asarray(Image.open(path.decode()))
        assert back.shape == (h, w, 4) and np.array_equal(back, px)


def test_rtrace_executable_fails_loudly_without_a_gpu():
    """The C++ host of the build (RTrace/main.swift over the C ABI): without a driver it reports and exits 1, no image."""
    import os
    import subprocess
    import torch
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpuraytracer_b200", "rtrace")
    assert os.path.exists(exe)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present: covered by the gpu tests")
    r = subprocess.run([exe, "/tmp/should_not_exist.png", "--spp", "1"], capture_output=True, text=True)
    assert r.returncode == 1 and "libcuda" in r.stderr and not os.path.exists("/tmp/should_not_exist.png")
