"""Known-answer tests that pin the oracle to values derived by hand from the
reference source; the reference itself has no tests or vectors."""
import ctypes as C
import math

import numpy as np

import oracle_lib as O


def f32(x):
    return np.float32(x)


def test_halton_known_values():
    # sampling.metal:107-122 evaluated by hand in fp32
    h = O.lib.rt_oracle_halton
    for d in range(24):
        assert h(0, d) == 0.0
    assert h(1, 0) == 0.5 and h(2, 0) == 0.25 and h(3, 0) == 0.75
    third = f32(1) / f32(3)
    assert f32(h(1, 1)) == f32(third * f32(1))
    ninth = f32(third * third)
    assert f32(h(5, 1)) == f32(f32(third * f32(2)) + f32(ninth * f32(1)))      # 5 = 12_3 -> 2/3 + 1/9


def test_halton_matches_python_loop():
    primes = [2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37, 41, 43, 47, 53, 59, 61, 67, 71, 73, 79, 83, 89]
    rng = np.random.default_rng(0)
    for i in rng.integers(0, 1 << 21, 200):
        for d in (0, 1, 2, 5, 10, 15, 23):
            b = primes[d]; f = f32(1); inv = f32(1) / f32(b); r = f32(0); k = int(i)
            while k > 0:
                f = f32(f * inv); r = f32(r + f32(f * f32(k % b))); k //= b
            assert f32(O.lib.rt_oracle_halton(int(i), d)) == r


def test_hash_vectors():
    def h(x):
        x ^= x >> 16; x = (x * 0x7feb352d) & 0xffffffff; x ^= x >> 15
        x = (x * 0x846ca68b) & 0xffffffff; x ^= x >> 16
        return x
    for x in (0, 1, 2, 12345, 0xdeadbeef, 0xffffffff):
        assert O.lib.rt_oracle_hash(x) == h(x)


def test_sincos_accuracy_and_quadrants():
    xs = np.linspace(0, 2 * math.pi, 20001).astype(np.float32)
    s = C.c_float(); c = C.c_float()
    worst = 0.0
    for x in xs[::7]:
        O.lib.rt_oracle_sincos(C.c_float(float(x)), C.byref(s), C.byref(c))
        worst = max(worst, abs(s.value - math.sin(float(x))), abs(c.value - math.cos(float(x))))
    assert worst < 2.5e-7


def test_cornell_fixture_shape():
    sc = O.cornell(800, 600)
    v = sc["vertices"].view(np.float32).reshape(108, 4)
    m = sc["materials"].view(np.float32).reshape(36, 12)
    assert sc["vertices"].nbytes == 1728 and sc["materials"].nbytes == 1728
    assert np.all(v[:, 3] == 0)
    # light appended last (ids 34, 35) at y = 2.49, emissive (1,1,1)   scene.swift:23-59
    assert np.all(v[102:, 1] == np.float32(2.5) - np.float32(0.01))
    assert np.all(m[34:, 8:11] == 1.0) and np.all(m[:34, 8:11] == 0.0)
    # back wall first triangle, scene.swift:81-85
    assert v[0].tolist() == [-2.5, -2.5, -2.5, 0] and v[1].tolist() == [2.5, 2.5, -2.5, 0]
    # red / green walls, scene.swift:72-73
    assert m[2, :3].tolist() == [np.float32(0.9), 0, 0] and m[4, :3].tolist() == [0, np.float32(0.7), 0]
    cam = sc["camera"].view(np.float32)
    assert cam[:3].tolist() == [0, 0, 9] and cam[4:7].tolist() == [0, 0, -1]
    assert sc["camera"].view(np.int32)[12:14].tolist() == [800, 600]


def test_center_pixel_zero_jitter_looks_down_minus_z():
    sc = O.cornell(512, 512)
    o = np.zeros(3, np.float32); d = np.zeros(3, np.float32)
    # pixel (256,256) with zero jitter: s = t = 0 -> dir = -w = (0,0,-1)
    O.lib.rt_oracle_camera_ray(O.vp(sc["camera"]), 256, 256, C.c_float(0), C.c_float(0), O.vp(o), O.vp(d))
    assert o.tolist() == [0, 0, 9] and d.tolist() == [0, 0, -1]


def test_area_light_sample_center():
    sc = O.cornell(512, 512)
    pos = np.array([0, 0, 0], np.float32); col = np.zeros(3, np.float32); ld = np.zeros(3, np.float32); dist = C.c_float()
    # u = (.5,.5) -> sample at light.center (sampling.metal:205-213); straight up from the origin
    O.lib.rt_oracle_sample_area_light(O.vp(sc["lights"]), C.c_float(.5), C.c_float(.5), O.vp(pos), O.vp(col), O.vp(ld), C.byref(dist))
    inv = np.float32(1) / np.float32(dist.value)
    # direction = (pos - P) * (1/dist): x*(1/x) may be one ulp below 1
    assert ld[0] == 0 and ld[2] == 0 and ld[1] == np.float32(dist.value) * inv and abs(dist.value - 2.49) < 1e-6
    assert np.allclose(col, np.array([1, .95, .9], np.float32) * inv * inv, rtol=1e-6)


def test_intersector_semantics():
    sc = O.cornell(512, 512)
    dist = C.c_float(); pid = C.c_uint32()
    o = np.array([0, 0, 9], np.float32); d = np.array([0, 0, -1], np.float32)
    hit = O.lib.rt_oracle_intersect(O.vp(sc["vertices"]), 36, None, 0, O.vp(o), O.vp(d), C.c_float(1e-3), C.c_float(1000), 0, C.byref(dist), C.byref(pid))
    # straight down the axis: the back wall diagonal is shared by triangles 0 and 1 -> lowest id wins
    assert hit == 1 and pid.value == 0 and dist.value == 11.5
    # tmax below the hit distance -> miss (strict t < tmax)
    hit = O.lib.rt_oracle_intersect(O.vp(sc["vertices"]), 36, None, 0, O.vp(o), O.vp(d), C.c_float(1e-3), C.c_float(11.5), 0, C.byref(dist), C.byref(pid))
    assert hit == 0
    # no culling: the same wall hit from behind, a little off the diagonal on either side of it
    for x in (0.25, -0.25):
        o2 = np.array([x, 0, -9], np.float32); d2 = np.array([0, 0, 1], np.float32)
        hit = O.lib.rt_oracle_intersect(O.vp(sc["vertices"]), 36, None, 0, O.vp(o2), O.vp(d2), C.c_float(1e-3), C.c_float(1000), 0, C.byref(dist), C.byref(pid))
        assert hit == 1 and pid.value in (0, 1) and dist.value == 6.5
    # the sign rule (rt_detmath.h): an exact zero of U or V is +0; from the front (D > 0) both triangles of the
    # shared diagonal accept it (checked above: id 0 wins), from behind neither does and the ray goes on
    o2 = np.array([0, 0, -9], np.float32); d2 = np.array([0, 0, 1], np.float32)
    hit = O.lib.rt_oracle_intersect(O.vp(sc["vertices"]), 36, None, 0, O.vp(o2), O.vp(d2), C.c_float(1e-3), C.c_float(1000), 0, C.byref(dist), C.byref(pid))
    assert hit == 0 or pid.value > 1


def test_image_statistics_and_determinism():
    sc = O.cornell(128, 96)
    a, cnt = O.pathtrace(sc, 128, 96, 8)
    b, _ = O.pathtrace(sc, 128, 96, 8, nthreads=1)
    assert np.array_equal(a, b)                       # thread count does not change a bit
    assert not np.isnan(a).any() and np.all(a[..., 3] == 1)
    black = (a[..., :3].sum(-1) == 0)
    # integer aspect ratio (sampling.metal:132) leaves a black border left and right of the box
    assert black[:, :3].all() and black[:, -3:].all() and 0.10 < black.mean() < 0.30
    assert int(cnt[0]) == 128 * 96 * 8 and cnt[1] >= cnt[0] and cnt[2] == cnt[3]


def test_tiles_and_progressive_are_bit_exact():
    sc = O.cornell(96, 64)
    full, _ = O.pathtrace(sc, 96, 64, 6)
    top, _ = O.pathtrace(sc, 96, 64, 6, row0=0, rows=20)
    bot, _ = O.pathtrace(sc, 96, 64, 6, row0=20, rows=44)
    assert np.array_equal(np.concatenate([top, bot]), full)
    acc = np.zeros((64, 96, 4), np.float32)
    for f in range(6):
        O.pathtrace(sc, 96, 64, 1, first_sample=f, accumulate=True, out=acc)
    assert np.all(acc[..., 3] == 6)
    assert np.array_equal(acc[..., :3] / np.float32(6), full[..., :3])


def test_bounces_limit():
    sc = O.cornell(16, 16)
    try:
        O.pathtrace(sc, 16, 16, 1, bounces=5)
        assert False
    except RuntimeError:
        pass


def test_light_extent_flag_samples_the_lights_own_rectangle():
    """RT_VARIANT_LIGHT_EXTENT (the TODO at sampling.metal:175 / the literals of :207-209): right and up
    span width x depth instead of 0.5 x 0.5.  By hand: a 1 x 2 light at (0, 2.49, 0), u = (1, 0) -> the corner
    (0.5, 2.49, -1); without the flag (0.25, 2.49, -0.25)."""
    light = O.cornell(64, 64)["lights"].copy()
    lf = light.view(np.float32)
    lf[0:3] = (0.0, 2.49, 0.0); lf[12], lf[13] = 1.0, 2.0
    col = np.zeros(3, np.float32); ld = np.zeros(3, np.float32); dist = C.c_float()
    for extent, corner in ((1, (0.5, 2.49, -1.0)), (0, (0.25, 2.49, -0.25))):
        pos = np.array([corner[0], 0.49, corner[2]], np.float32)          # two units straight below the sample
        O.lib.rt_oracle_sample_area_light_ex(O.vp(light), C.c_float(1.0), C.c_float(0.0), O.vp(pos), O.vp(col), O.vp(ld),
                                             C.byref(dist), extent)
        assert ld.tolist() == [0.0, 1.0, 0.0] and dist.value == np.float32(2.49) - np.float32(0.49)
        inv = np.float32(1) / np.float32(dist.value)
        assert np.array_equal(col, lf[4:7] * np.float32(inv * inv))
    # whole frames: a 0.5 x 0.5 light under the flag is the reference's literal (0.5 * 0.5 = 0.25 exactly) ...
    W, H, spp = 64, 48, 4
    sc = O.cornell(W, H)
    ref, _ = O.pathtrace(sc, W, H, spp)
    assert sc["lights"].view(np.float32)[12] == 1.0 and sc["lights"].view(np.float32)[13] == 1.0    # scene.swift:23-34
    wide, cnt_wide = O.pathtrace(sc, W, H, spp, light_extent=True)
    assert not np.array_equal(wide, ref)                                  # ... the Cornell light is 1 x 1: other samples
    sc["lights"].view(np.float32)[12:14] = 0.5
    same, _ = O.pathtrace(sc, W, H, spp, light_extent=True)
    assert np.array_equal(same, ref)


def test_half_resolve_is_ieee_round_to_nearest_even():
    """rt_oracle_resolve_rgba16f (the checker of rtResolveToRGBA16F; the reference's render target is rgba16Float,
    renderer.swift:74-82): by hand 1 + 2^-11 ties to 1.0 (0x3C00), 1 + 3 * 2^-11 ties to 1 + 2^-9 (0x3C02), 65520 -> inf,
    2^-25 -> 0, anything above it -> the smallest subnormal; and numpy's float16 cast agrees on 2^20 random bit patterns."""
    def half(values, divide=0):
        f = np.asarray(values, np.float32).reshape(-1, 4).copy()
        out = np.zeros(f.shape, np.uint16)
        O.lib.rt_oracle_resolve_rgba16f(O.vp(f), O.vp(out), C.c_uint64(len(f)), divide)
        return out
    e = half([1.0 + 2.0 ** -11, 1.0 + 3 * 2.0 ** -11, 65520.0, 2.0 ** -25,
              np.float32(2.0 ** -25) * np.float32(1.0000001), -2.0, 65504.0, 0.0])
    assert e.reshape(-1).tolist() == [0x3C00, 0x3C02, 0x7C00, 0x0000, 0x0001, 0xC000, 0x7BFF, 0x0000]
    bits = np.random.default_rng(7).integers(0, 1 << 32, 1 << 20, dtype=np.uint64).astype(np.uint32)
    f = bits.view(np.float32)
    with np.errstate(all="ignore"):
        want = f.astype(np.float16).view(np.uint16)
    got = half(f).reshape(-1)
    ok = ~np.isnan(f)
    assert np.array_equal(got[ok], want[ok]) and np.all(np.isnan(got[~ok].view(np.float16)))
    # progressive buffers: xyz / w, alpha 1
    assert half([3.0, 6.0, 1.5, 3.0], 1).reshape(-1).tolist() == [0x3C00, 0x4000, 0x3800, 0x3C00]
