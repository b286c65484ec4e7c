"""GPU parity: the CUDA path, called through the C ABI (librtcuda.so), against the
oracle on the same seeded inputs.  Tolerance: north_star asks for 1e-4 relative per
pixel/channel in fp32; because both sides follow the same arithmetic contract the
tests first demand bit equality and report the 1e-4 figure beside it."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as O
import gpuraytracer_b200 as G
from gpuraytracer_b200 import _lib as L

pytestmark = pytest.mark.gpu

REL_TOL = 1e-4   # north_star: "within 1e-4 relative for fp32"


def out_of_tolerance(gpu, ref):
    err = np.abs(gpu - ref)
    return int((err > REL_TOL * np.maximum(np.abs(ref), 1e-6)).sum())


@pytest.fixture(scope="module")
def dev():
    return G.Device(0)


def debug_eval(dev, what, inp, out):
    L.check(L.lib.rtDebugEvaluate(dev.h, what, inp.ctypes.data, inp.nbytes, out.ctypes.data, out.nbytes, out.shape[0]),
            "rtDebugEvaluate")
    return out


def test_device_is_b200(dev):
    i = dev.info()
    assert i.ccMajor == 10 and i.smCount >= 100 and i.supportsRaytracing == 1


def test_halton_bit_exact(dev):
    rng = np.random.default_rng(1)
    n = 1 << 16
    idx = rng.integers(0, (1 << 20) + 4096, n).astype(np.uint32)
    idx[:8] = [0, 1, 2, 3, (1 << 22) - 1, 1 << 22, (1 << 24) + 5, 0xffffffff]    # beyond the unrolled digits too
    dim = rng.integers(0, 24, n).astype(np.uint32)
    pairs = np.stack([idx, dim], 1).copy()
    ref = np.array([O.lib.rt_oracle_halton(int(i), int(d)) for i, d in pairs], np.float32)
    for what in (0, 3):
        got = debug_eval(dev, what, pairs, np.zeros(n, np.float32))
        assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), what


def test_halton_slot_table_form_bit_exact(dev):
    """halton_slot<s>(i, bounce): the table-driven form the rolled bounce loop uses (dimension 2+5b+s)."""
    rng = np.random.default_rng(5)
    n = 1 << 16
    idx = rng.integers(0, (1 << 20) + 4096, n).astype(np.uint32)
    idx[:8] = [0, 1, 2, 3, (1 << 22) - 1, 1 << 22, (1 << 24) + 5, 0xffffffff]
    code = rng.integers(0, 16, n).astype(np.uint32)              # 4*bounce + slot
    pairs = np.stack([idx, code], 1).copy()
    ref = np.array([O.lib.rt_oracle_halton(int(i), 2 + 5 * (int(c) >> 2) + (int(c) & 3)) for i, c in pairs], np.float32)
    got = debug_eval(dev, 4, pairs, np.zeros(n, np.float32))
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))


def test_sincos_and_hash_bit_exact(dev):
    x = np.concatenate([np.linspace(0, 6.2831855, 100000), np.random.default_rng(2).uniform(-50, 50, 31072)]).astype(np.float32)
    got = debug_eval(dev, 1, x, np.zeros((x.size, 2), np.float32))
    s = C.c_float(); c = C.c_float()
    ref = np.zeros_like(got)
    for k, v in enumerate(x):
        O.lib.rt_oracle_sincos(C.c_float(float(v)), C.byref(s), C.byref(c))
        ref[k] = (s.value, c.value)
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))
    u = np.arange(0, 1 << 16, dtype=np.uint32) * np.uint32(2654435761)
    got = debug_eval(dev, 2, u, np.zeros(u.size, np.uint32))
    assert np.array_equal(got, np.array([O.lib.rt_oracle_hash(int(v)) for v in u], np.uint32))


@pytest.mark.parametrize("variant", [L.VARIANT_SMEM, L.VARIANT_CONST, L.VARIANT_WAVE, L.VARIANT_SORTED, L.VARIANT_SORTED_BLOCK])
def test_config1_cornell_512_16spp(dev, variant):
    """BASELINE.json configs[0]: Cornell 512x512, 16 spp."""
    W = H = 512
    ref, _ = O.pathtrace(O.cornell(W, H), W, H, 16)
    r = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev)
    img = r.draw(spp=16, variant=variant)
    assert out_of_tolerance(img, ref) == 0
    assert np.array_equal(img, ref), f"{(img != ref).sum()} values differ"
    assert r.last_command.launch_count() == 2 and r.last_command.gpu_time_ms() > 0


def test_reference_literals_400spp_3bounces(dev):
    """No rtSetParams-style overrides beyond the grid: the shader's own 400 spp / 3 bounces."""
    W, H = 96, 72
    ref, _ = O.pathtrace(O.cornell(W, H), W, H, 400)
    img = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev).draw()
    assert np.array_equal(img, ref)


@pytest.mark.parametrize("bounces", [1, 2, 4])
def test_bounce_counts(dev, bounces):
    W, H = 160, 120
    ref, _ = O.pathtrace(O.cornell(W, H), W, H, 8, bounces=bounces)
    img = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev).draw(spp=8, bounces=bounces)
    assert np.array_equal(img, ref)


def test_ragged_sizes_and_row_tiles(dev):
    """Sizes that are not multiples of the 16x8 block, rendered whole and as row tiles."""
    W, H = 203, 77
    sc = G.init_cornell_box(W, H)
    ref, _ = O.pathtrace(O.cornell(W, H), W, H, 4)
    assert np.array_equal(G.Renderer(sc, W, H, device=dev).draw(spp=4), ref)
    parts = [G.Renderer(sc, W, H, device=dev, row0=r0, rows=n).draw(spp=4) for r0, n in ((0, 13), (13, 1), (14, 63))]
    assert np.array_equal(np.concatenate(parts), ref)


def test_progressive_accumulation_is_bit_exact(dev):
    """Config 5 in miniature: 1 spp per frame, accumulate, resolve == one launch."""
    W, H, N = 128, 96, 12
    ref, _ = O.pathtrace(O.cornell(W, H), W, H, N)
    r = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev)
    r.clear()
    for f in range(N):
        r.draw_async(spp=1, firstSample=f, accumulate=1)
    r.last_command.wait_until_completed()
    acc = r.image()
    assert np.all(acc[..., 3] == N)
    assert np.array_equal(acc[..., :3] / np.float32(N), ref[..., :3])


def test_spheres_scene_config4_small(dev):
    W, H = 192, 108
    ref, cnt = O.pathtrace(O.spheres(W, H), W, H, 4)
    for variant in (L.VARIANT_SMEM, L.VARIANT_CONST, L.VARIANT_WAVE):
        img = G.Renderer(G.init_sphere_scene(W, H), W, H, device=dev).draw(spp=4, variant=variant)
        assert out_of_tolerance(img, ref) == 0
        assert np.array_equal(img, ref)
    assert cnt[2] == 4 * cnt[3]          # four lights -> four shadow rays per shaded bounce


def test_1080p_subset_rows_full_width(dev):
    """Bench shape (1920x1080) on a row subset the oracle finishes in seconds."""
    W, H = 1920, 1080
    sc = G.init_cornell_box(W, H)
    for row0 in (0, 536, 1072):
        ref, _ = O.pathtrace(O.cornell(W, H), W, H, 8, row0=row0, rows=8)
        img = G.Renderer(sc, W, H, device=dev, row0=row0, rows=8).draw(spp=8)
        assert np.array_equal(img, ref), row0


def test_resolve_rgba8_matches_image_swift(dev):
    W, H = 128, 96
    r = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev)
    img = r.draw(spp=16)
    want = np.zeros((H, W, 4), np.uint8)
    O.lib.rt_oracle_tonemap_rgba8(O.vp(img), O.vp(want), C.c_uint64(W * H))
    got = r.resolve_rgba8()
    assert np.abs(got.astype(int) - want.astype(int)).max() <= 1 and np.all(got[..., 3] == 255)


def test_resolve_rgba16f_is_the_reference_render_target_format(dev):
    """rtResolveToRGBA16F: the fp32 image as the rgba16Float texture the reference renders into (renderer.swift:74-82),
    every channel rounded to nearest-even -- the [Float16] array image.swift:25-38 reads.  Bit-equal to the oracle's
    conversion, also from a progressive accumulation buffer (xyz / w, alpha 1) and for values that overflow,
    underflow and tie."""
    W, H = 128, 96
    r = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev)
    img = r.draw(spp=16)
    want = np.zeros((H, W, 4), np.uint16)
    O.lib.rt_oracle_resolve_rgba16f(O.vp(img), O.vp(want), C.c_uint64(W * H), 0)
    got = r.resolve_rgba16f()
    assert np.array_equal(got.view(np.uint16), want) and np.all(got[..., 3] == 1.0)
    r.clear()
    r.draw_async(spp=5, firstSample=0, accumulate=1)
    r.draw_async(spp=3, firstSample=5, accumulate=1).wait_until_completed()
    acc = r.image()
    O.lib.rt_oracle_resolve_rgba16f(O.vp(acc), O.vp(want), C.c_uint64(W * H), 1)
    assert np.array_equal(r.resolve_rgba16f(divide_by_alpha=True).view(np.uint16), want)
    # arbitrary bit patterns through the same pass (no NaNs: their payload is not part of the contract)
    bits = np.random.default_rng(3).integers(0, 1 << 32, W * H * 4, dtype=np.uint64).astype(np.uint32)
    vals = bits.view(np.float32).copy()
    vals[np.isnan(vals)] = 65520.0
    vals[:8] = (1.0 + 2.0 ** -11, 1.0 + 3 * 2.0 ** -11, 65520.0, 2.0 ** -25, 6e-8, -2.0, 65504.0, 0.0)
    r.render_buffer.contents(np.float32)[:] = vals
    r.render_buffer.did_modify()
    O.lib.rt_oracle_resolve_rgba16f(O.vp(vals), O.vp(want), C.c_uint64(W * H), 0)
    assert np.array_equal(r.resolve_rgba16f().view(np.uint16), want)


def test_error_paths(dev):
    with pytest.raises(G.RTError) as e:
        dev.make_pipeline("noSuchKernel")
    assert e.value.code == -2
    r = G.Renderer(G.init_cornell_box(64, 64), 64, 64, device=dev)
    with pytest.raises(G.RTError) as e:
        r.draw(bounces=5)
    assert e.value.code == -7
    cmd = G.Command(r.queue, r.pipeline)
    cmd.dispatch_threads((64, 64, 1))
    with pytest.raises(G.RTError) as e:
        cmd.commit()
    assert e.value.code == -4                                   # nothing bound
    with pytest.raises(G.RTError) as e:
        G.Renderer(G.init_cornell_box(64, 64), 64, 64, device=dev, rows=32).encode(rows=64).commit()
    assert e.value.code in (-1, -5)


def test_interleaved_tiles_and_exact_counters(dev):
    """The multi-GPU deal on one GPU: every rank's interleaved rows, stitched, equal the full frame;
    the counting variant reproduces the oracle's integer work counters exactly."""
    from gpuraytracer_b200 import distributed as D
    W, H, spp, world, tile = 208, 100, 4, 3, 8
    sc = G.init_cornell_box(W, H)
    ref, cnt = O.pathtrace(O.cornell(W, H), W, H, spp)
    off_full = G.fill_offsets(W, 0, H).reshape(H, W)
    idx, pad = D.scatter_index(H, world, tile)
    parts = []
    for rnk in range(world):
        row0, rows, tr, ts = D.tile_params(H, rnk, world, tile)
        mine = D.local_rows(H, rnk, world, tile)
        r = G.Renderer(sc, W, H, device=dev, row0=row0, rows=rows, offsets=np.ascontiguousarray(off_full[mine]).reshape(-1))
        img = r.draw(spp=spp, tileRows=tr, tileStride=ts)
        parts.append(np.concatenate([img, np.zeros((pad - rows, W, 4), np.float32)]))
    assert np.array_equal(np.concatenate(parts)[idx], ref)
    r = G.Renderer(sc, W, H, device=dev)
    for variant in (L.VARIANT_SMEM, L.VARIANT_WAVE):
        c = r.count_rays(spp=spp, variant=variant)
        assert [c["paths"], c["closest"], c["shadow"], c["shaded"]] == [int(v) for v in cnt[:4]], variant
        assert np.array_equal(r.image(), ref)


def test_wave_kernel_odd_sample_counts_and_sizes(dev):
    """The compacting kernel traces 8 samples per pixel per round: sample counts that are not multiples
    of 8, ragged image sizes (lanes without a pixel), accumulation across launches, 1..4 bounces."""
    W, H = 203, 77
    sc, osc = G.init_cornell_box(W, H), O.cornell(W, H)
    for spp, bounces in ((5, 3), (13, 2), (8, 4), (9, 1)):
        ref, _ = O.pathtrace(osc, W, H, spp, bounces=bounces)
        img = G.Renderer(sc, W, H, device=dev).draw(spp=spp, bounces=bounces, variant=L.VARIANT_WAVE)
        assert np.array_equal(img, ref), (spp, bounces)
    ref, _ = O.pathtrace(osc, W, H, 20)
    r = G.Renderer(sc, W, H, device=dev)
    r.clear()
    r.draw_async(spp=7, firstSample=0, accumulate=1, variant=L.VARIANT_WAVE)
    r.draw_async(spp=13, firstSample=7, accumulate=1, variant=L.VARIANT_WAVE).wait_until_completed()
    acc = r.image()
    assert np.all(acc[..., 3] == 20) and np.array_equal(acc[..., :3] / np.float32(20), ref[..., :3])


def test_group_render_single_process_row_tiles(dev):
    """rtGroupRender: one process, interleaved 8-row tiles, every member's tiles copied device to device into place in
    member 0's frame, one read-back.  Several group members on the same GPU exercise the deal, the strided copies and the
    short last tile (90 rows = 11 tiles + 2 rows) without needing a second device; tools/gpu_scale.sh runs it on 2-8."""
    W, H, spp = 160, 90, 6
    osc = O.cornell(W, H)
    ref, _ = O.pathtrace(osc, W, H, spp)
    off = O.offsets(W, 0, H)
    for ordinals in ([0], [0, 0], [0, 0, 0], [0] * 5):
        arr = (C.c_int * len(ordinals))(*ordinals)
        g = C.c_void_p()
        L.check(L.lib.rtGroupCreate(arr, len(ordinals), C.byref(g)), "rtGroupCreate")
        assert L.lib.rtGroupDeviceCount(g) == len(ordinals)
        ms = C.c_double()

        def render(sc, p, out, W=W, H=H, off=off):
            L.check(L.lib.rtGroupRender(g, b"pathTrace", O.vp(sc["camera"]), O.vp(sc["materials"]), sc["materials"].nbytes,
                                        O.vp(sc["lights"]), sc["lights"].nbytes, O.vp(sc["vertices"]), sc["vertices"].nbytes,
                                        O.vp(sc["spheres"]), 0 if sc["spheres"] is None else sc["spheres"].nbytes,
                                        O.vp(off), C.byref(p), O.vp(out), C.byref(ms)), "rtGroupRender")
        out = np.zeros((H, W, 4), np.float32)
        for call in range(2):                                       # the second call reuses every object of the first
            out[:] = -1
            render(osc, G.default_params(W, H, spp=spp, numTriangles=36, numLights=1), out)
            assert ms.value > 0 and np.array_equal(out, ref), (ordinals, call)
        # progressive: 2 + 4 samples accumulated through hostOut
        acc = np.zeros((H, W, 4), np.float32)
        render(osc, G.default_params(W, H, spp=2, firstSample=0, accumulate=1, numTriangles=36, numLights=1), acc)
        render(osc, G.default_params(W, H, spp=4, firstSample=2, accumulate=1, numTriangles=36, numLights=1), acc)
        assert np.all(acc[..., 3] == spp) and np.array_equal(acc[..., :3] / np.float32(spp), ref[..., :3]), ordinals
        # another scene, another size, through the same group: spheres + four lights, 13 rows short of a tile
        W2, H2 = 96, 77
        ssc = O.spheres(W2, H2)
        ref2, _ = O.pathtrace(ssc, W2, H2, 3)
        out2 = np.zeros((H2, W2, 4), np.float32)
        render(ssc, G.default_params(W2, H2, spp=3, numTriangles=ssc["T"], numSpheres=ssc["S"], numLights=ssc["L"]), out2,
               W=W2, H=H2, off=O.offsets(W2, 0, H2))
        assert np.array_equal(out2, ref2), ordinals
        L.check(L.lib.rtGroupRelease(g), "rtGroupRelease")


def _set_camera(cam_bytes, pos, direction, up=(0.0, 1.0, 0.0), fov=None):
    f = cam_bytes.view(np.float32)               # CameraGPU: position@0, direction@16, up@32, fov@56 (shaderTypes.h)
    f[0:3] = pos; f[4:7] = direction; f[8:11] = up
    if fov is not None:
        f[14] = fov


@pytest.mark.parametrize("scene_name", ["cornell", "spheres"])
def test_camera_candidate_lists_moved_cameras(dev, scene_name):
    """The wave kernel tests camera rays only against the primitives whose projection touches the warp's
    pixels (camera_candidates).  Cameras that put primitives behind the eye, edge-on (eye in a wall's plane),
    across the whole screen (eye inside the sphere lattice), rolled and wide-angle must still give the
    oracle's image bit for bit."""
    W, H, spp = 96, 64, 4
    cams = [((0.0, 0.0, 9.0), (0.0, 0.0, -1.0), (0.0, 1.0, 0.0), None),          # the fixture's own
            ((0.3, -0.4, 1.0), (0.2, -0.1, -1.0), (0.0, 1.0, 0.0), 1.6),           # inside the room, walls behind the eye
            ((2.5, 0.0, 2.0), (-0.3, 0.0, -1.0), (0.0, 1.0, 0.0), 1.2),            # eye in the plane of the right wall
            ((0.0, -2.5, 6.0), (0.0, 0.02, -1.0), (0.0, 1.0, 0.0), None),          # eye in the floor plane
            ((-1.2, 1.9, 0.4), (0.5, -0.8, -0.3), (0.3, 1.0, 0.1), 2.4),           # looking down, rolled, very wide
            ((0.0, 0.0, -1.0), (0.0, 0.3, 1.0), (0.0, 1.0, 0.0), 1.0)]             # looking out of the open front
    for pos, direction, up, fov in cams:
        osc = O.cornell(W, H) if scene_name == "cornell" else O.spheres(W, H)
        gsc = G.init_cornell_box(W, H) if scene_name == "cornell" else G.init_sphere_scene(W, H)
        _set_camera(osc["camera"], pos, direction, up, fov)
        _set_camera(gsc.camera, pos, direction, up, fov)
        assert np.array_equal(osc["camera"], gsc.camera)
        ref, _ = O.pathtrace(osc, W, H, spp)
        r = G.Renderer(gsc, W, H, device=dev)
        img = r.draw(spp=spp, variant=L.VARIANT_WAVE)
        full = r.draw(spp=spp, variant=L.VARIANT_SMEM)                    # the kernel without candidate lists
        assert np.array_equal(full, ref), (pos, "smem")
        assert np.array_equal(img, ref), (pos, "wave", int((img != ref).sum()))


def test_persistent_tiles_with_sample_chunks(dev):
    """pathTrace_tiles: (8x4 tile, sample chunk) units from an atomic counter; chunk c continues the running sum of
    chunk c-1 through the output buffer.  Same bits as one pass, whatever the chunking, also with row tiles,
    ragged sizes, accumulation onto an existing image, and as the library's own choice for a small share."""
    W, H, spp = 200, 77, 72
    osc = O.cornell(W, H)
    ref, _ = O.pathtrace(osc, W, H, spp)
    r = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev)
    for chunks in ("1", "2", "3", "9"):
        dev.set_tuning("tiles_chunks", int(chunks))
        img = r.draw(spp=spp, variant=L.VARIANT_TILES)
        assert np.array_equal(img, ref), chunks
    dev.set_tuning("tiles_chunks", 0)
    assert np.array_equal(r.draw(spp=spp, variant=L.VARIANT_TILES), ref)
    assert np.array_equal(r.draw(spp=spp), ref)                      # default: few blocks -> the library picks tiles
    # progressive: 40 + 32 samples accumulated, second launch chunked
    dev.set_tuning("tiles_chunks", 2)
    r.clear()
    r.draw_async(spp=40, firstSample=0, accumulate=1, variant=L.VARIANT_TILES)
    r.draw_async(spp=32, firstSample=40, accumulate=1, variant=L.VARIANT_TILES).wait_until_completed()
    acc = r.image()
    assert np.all(acc[..., 3] == spp) and np.array_equal(acc[..., :3] / np.float32(spp), ref[..., :3])
    # interleaved row tiles of rank 1 of 3
    from gpuraytracer_b200 import distributed as D
    row0, rows, tr, ts = D.tile_params(H, 1, 3, 8)
    mine = D.local_rows(H, 1, 3, 8)
    off = np.ascontiguousarray(G.fill_offsets(W, 0, H, seed=1).reshape(H, W)[mine]).reshape(-1)
    r2 = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev, row0=row0, rows=rows, offsets=off)
    assert np.array_equal(r2.draw(spp=spp, variant=L.VARIANT_TILES, tileRows=tr, tileStride=ts), ref[mine])
    dev.set_tuning("tiles_chunks", 0)


def test_triangle_test_exact_zero_cases(dev):
    """The sign rule makes the sign bit of an exact zero of U, V, Nn or D decide (rays through edges and vertices,
    origins in the triangle's plane, rays in the plane).  Small-integer triangles and rays produce such zeros in
    bulk: the device test must agree with the oracle's on every one, bit for bit (a compiler that moves a negation
    through the arithmetic breaks exactly these -- DESIGN.md 3)."""
    rng = np.random.default_rng(7)
    n = 60000
    cases = np.zeros((n, 17), np.float32)
    cases[:, 0:9] = rng.integers(-3, 4, (n, 9))                  # v0 v1 v2 on the integer lattice
    cases[:, 9:12] = rng.integers(-4, 5, (n, 3))                 # origin
    tgt = rng.integers(-3, 4, (n, 3)).astype(np.float32)         # aim at lattice points: edges, vertices, planes
    half = rng.random(n) < 0.5
    tgt[half] = (cases[half, 0:3] + cases[half, 3:6]) * np.float32(0.5)   # ... and at edge midpoints
    cases[:, 12:15] = tgt - cases[:, 9:12]
    cases[:, 15] = np.where(rng.random(n) < 0.5, 0.0, 1e-3)      # shadow-ray and camera-ray tmin
    cases[:, 16] = np.where(rng.random(n) < 0.3, 1.0, 1000.0)    # tmax == the hit distance for many
    out = np.zeros((n, 3), np.uint32)
    debug_eval(dev, 5, cases.reshape(-1), out)
    hits = 0
    dist = C.c_float(); pid = C.c_uint32()
    for k in range(n):
        verts = np.zeros((3, 4), np.float32); verts[:, :3] = cases[k, 0:9].reshape(3, 3)
        o = np.ascontiguousarray(cases[k, 9:12]); d = np.ascontiguousarray(cases[k, 12:15])
        hit = O.lib.rt_oracle_intersect(O.vp(verts), 1, None, 0, O.vp(o), O.vp(d), C.c_float(float(cases[k, 15])),
                                        C.c_float(float(cases[k, 16])), 0, C.byref(dist), C.byref(pid))
        assert hit == int(out[k, 0]), (k, cases[k].tolist())
        if hit:
            hits += 1
            assert np.float32(dist.value).view(np.uint32) == out[k, 1], (k, cases[k].tolist())
    assert hits > n // 20            # the set does exercise accepted hits, not only rejections


def test_rtrace_executable_writes_the_oracles_image(dev, tmp_path):
    """gpuraytracer_b200/rtrace = RTrace/main.swift over the C ABI (csrc/host_renderer.cpp mirrors renderer.swift and
    image.swift): its PNG must be the oracle's image through the oracle's tone map, +-1 LSB (powf on the device)."""
    import os
    import subprocess
    from PIL import Image
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpuraytracer_b200", "rtrace")
    W, H, spp = 160, 120, 16
    out = str(tmp_path / "out.png")
    r = subprocess.run([exe, out, "--width", str(W), "--height", str(H), "--spp", str(spp)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "Image saved to" in r.stdout and "Render completed in" in r.stdout
    png = np.asarray(Image.open(out))
    ref, _ = O.pathtrace(O.cornell(W, H), W, H, spp)
    want = np.zeros((H, W, 4), np.uint8)
    O.lib.rt_oracle_tonemap_rgba8(O.vp(np.ascontiguousarray(ref)), O.vp(want), C.c_uint64(W * H))
    assert png.shape == (H, W, 4) and np.all(png[..., 3] == 255)
    assert int(np.abs(png[..., :3].astype(int) - want[..., :3].astype(int)).max()) <= 1
    # unknown option -> usage error, no crash
    assert subprocess.run([exe, "--nonsense"], capture_output=True).returncode == 2


def test_wide_kernel_few_samples_per_launch(dev):
    """pathTrace_wide (launches with fewer samples than a round has slots): a warp takes 8 / spp tiles side by side.
    Same bits as the oracle for every spp < 8, ragged widths, row tiles, and as a progressive accumulation of
    1-spp frames (the library's default for them)."""
    W, H = 203, 61                      # 203 = 25 tiles + 3 pixels: ragged against every G
    osc = O.cornell(W, H)
    r = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev)
    for spp in (1, 2, 3, 4, 5, 7):
        ref, _ = O.pathtrace(osc, W, H, spp)
        assert np.array_equal(r.draw(spp=spp, variant=L.VARIANT_WIDE), ref), spp
        assert np.array_equal(r.draw(spp=spp), ref), spp           # default picks the wide form below 8 spp
    # progressive: 12 frames of 1 spp == one 12-spp launch, bit for bit
    ref, _ = O.pathtrace(osc, W, H, 12)
    r.clear()
    for f in range(12):
        cmd = r.draw_async(spp=1, firstSample=f, accumulate=1)
    cmd.wait_until_completed()
    acc = r.image()
    assert np.all(acc[..., 3] == 12) and np.array_equal(acc[..., :3] / np.float32(12), ref[..., :3])
    # spheres scene (R = 4 there: wide4) and a row tile
    W2, H2 = 120, 40
    ref, _ = O.pathtrace(O.spheres(W2, H2), W2, H2, 2, row0=8, rows=24)
    off = G.fill_offsets(W2, 8, 24, seed=1)
    r2 = G.Renderer(G.init_sphere_scene(W2, H2), W2, H2, device=dev, row0=8, rows=24, offsets=off)
    assert np.array_equal(r2.draw(spp=2), ref)


def test_paired_and_unpaired_triangle_loops_agree(dev):
    """rt_prep_scene flags a scene whose triangles 2k, 2k+1 share v0 bit for bit (quads cut in two); the wave kernel
    then reuses w and cross(d, w) for the second triangle of each pair.  Breaking the pairing (one vertex moved by
    one ulp, or an odd triangle count) must switch to the plain loops, and both must match the oracle."""
    W, H, spp = 128, 96, 8
    for mode in ("paired", "ulp", "odd"):
        osc = O.cornell(W, H); gsc = G.init_cornell_box(W, H)
        T = 36
        if mode == "ulp":
            for sc_bytes in (osc["vertices"], gsc.vertices):
                v = sc_bytes.view(np.uint32)
                v[3 * 4 * 5 + 0] += 1            # triangle 5, vertex 0, x: one ulp off its partner's v0
        if mode == "odd":
            T = 35                               # drop the last light triangle: T odd -> not paired
            osc["vertices"] = osc["vertices"][:T * 48].copy(); osc["materials"] = osc["materials"][:T * 48].copy(); osc["T"] = T
            gsc.vertices = gsc.vertices[:T * 48].copy(); gsc.materials = gsc.materials[:T * 48].copy(); gsc.num_triangles = T
        ref, _ = O.pathtrace(osc, W, H, spp)
        r = G.Renderer(gsc, W, H, device=dev)
        assert np.array_equal(r.draw(spp=spp, variant=L.VARIANT_WAVE), ref), mode
        assert np.array_equal(r.draw(spp=spp, variant=L.VARIANT_SMEM), ref), mode


def _quad(v, t, a, b, c, d):
    """triangles t, t+1 = (a, b, c), (a, c, d): a quad cut in two on a shared v0, as scene.swift builds them"""
    f = v.view(np.float32)
    for k, p in enumerate((a, b, c, a, c, d)):
        f[(3 * t + k) * 4:(3 * t + k) * 4 + 3] = p


@pytest.mark.parametrize("variant", [L.VARIANT_SORTED, L.VARIANT_SORTED_BLOCK])
def test_sorted_kernel_moved_cameras_bounces_and_sizes(dev, variant):
    """pathTrace_sorted (csrc/sorted.cuh; sortedblk.cuh: the sort taken over the block): the paths of a round sorted
    by the object groups their rays can reach.
    Same bits as the oracle for cameras inside the room / in a wall's plane, every bounce count, odd sample counts
    and ragged sizes, progressive accumulation, and in its ticketed form with sample chunks."""
    W, H = 96, 64
    cams = [((0.0, 0.0, 9.0), (0.0, 0.0, -1.0), (0.0, 1.0, 0.0), None),
            ((0.3, -0.4, 1.0), (0.2, -0.1, -1.0), (0.0, 1.0, 0.0), 1.6),
            ((2.5, 0.0, 2.0), (-0.3, 0.0, -1.0), (0.0, 1.0, 0.0), 1.2),
            ((-1.2, 1.9, 0.4), (0.5, -0.8, -0.3), (0.3, 1.0, 0.1), 2.4)]
    for (pos, direction, up, fov), bounces, spp in zip(cams, (3, 4, 2, 1), (12, 7, 13, 6)):
        osc = O.cornell(W, H); gsc = G.init_cornell_box(W, H)
        _set_camera(osc["camera"], pos, direction, up, fov)
        _set_camera(gsc.camera, pos, direction, up, fov)
        ref, _ = O.pathtrace(osc, W, H, spp, bounces=bounces)
        img = G.Renderer(gsc, W, H, device=dev).draw(spp=spp, bounces=bounces, variant=variant)
        assert np.array_equal(img, ref), (pos, bounces, int((img != ref).sum()))
    W, H, spp = 203, 75, 72
    ref, _ = O.pathtrace(O.cornell(W, H), W, H, spp)
    r = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev)
    assert np.array_equal(r.draw(spp=spp, variant=variant), ref)
    r.clear()
    r.draw_async(spp=41, firstSample=0, accumulate=1, variant=variant)
    r.draw_async(spp=31, firstSample=41, accumulate=1, variant=variant).wait_until_completed()
    acc = r.image()
    assert np.all(acc[..., 3] == spp) and np.array_equal(acc[..., :3] / np.float32(spp), ref[..., :3])


@pytest.mark.parametrize("variant", [L.VARIANT_SORTED_TILES, L.VARIANT_SORTED_BLOCK_TILES])
def test_sorted_tiles_with_sample_chunks(dev, variant):
    W, H, spp = 200, 77, 72
    ref, _ = O.pathtrace(O.cornell(W, H), W, H, spp)
    r = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev)
    for chunks in ("1", "2", "3", "9"):
        dev.set_tuning("tiles_chunks", int(chunks))
        assert np.array_equal(r.draw(spp=spp, variant=variant), ref), chunks
    dev.set_tuning("tiles_chunks", 0)
    assert np.array_equal(r.draw(spp=spp), ref)                      # default: few blocks -> the ticketed sorted form
    from gpuraytracer_b200 import distributed as D
    row0, rows, tr, ts = D.tile_params(H, 1, 3, 8)
    mine = D.local_rows(H, 1, 3, 8)
    off = np.ascontiguousarray(G.fill_offsets(W, 0, H, seed=1).reshape(H, W)[mine]).reshape(-1)
    r2 = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev, row0=row0, rows=rows, offsets=off)
    assert np.array_equal(r2.draw(spp=spp, variant=variant, tileRows=tr, tileStride=ts), ref[mine])


def test_sorted_kernel_other_groupings(dev):
    """rt_prep_scene's object groups are whatever the vertices say: scenes whose quads fall into other groups than
    the Cornell ones (free-floating quads: more groups than key bits; a box welded to the room: fewer; nothing
    cullable or pairs broken: the kernel's own fall-back to the compacting form) all give the oracle's image."""
    W, H, spp = 112, 80, 12
    rng = np.random.default_rng(7)
    for mode in ("floating", "clusters", "welded", "room_only", "ulp"):
        osc = O.cornell(W, H); gsc = G.init_cornell_box(W, H)
        T = 36
        for sc_v in (osc["vertices"], gsc.vertices):
            v = sc_v
            g = np.random.default_rng(11)
            if mode == "floating":            # the 12 box quads become free quads anywhere in the room
                for q in range(12):
                    a = g.uniform(-2.2, 2.2, 3).astype(np.float32)
                    e1 = g.uniform(-0.9, 0.9, 3).astype(np.float32); e2 = g.uniform(-0.9, 0.9, 3).astype(np.float32)
                    _quad(v, 10 + 2 * q, a, a + e1, a + e1 + e2, a + e2)
            elif mode == "clusters":          # three fans of four quads, each fan on one shared corner
                for c in range(3):
                    a = g.uniform(-1.8, 1.8, 3).astype(np.float32)
                    for q in range(4):
                        e1 = g.uniform(-0.8, 0.8, 3).astype(np.float32); e2 = g.uniform(-0.8, 0.8, 3).astype(np.float32)
                        _quad(v, 10 + 2 * (4 * c + q), a, a + e1, a + e1 + e2, a + e2)
            elif mode == "welded":            # one quad of the tall box gets the room's corner as a vertex
                f = v.view(np.float32)
                corner = np.array([-2.5, -2.5, -2.5], np.float32)
                a = f[(3 * 10) * 4:(3 * 10) * 4 + 3].copy(); b = f[(3 * 10 + 1) * 4:(3 * 10 + 1) * 4 + 3].copy()
                _quad(v, 10, a, b, corner, corner + np.float32(0.5) * (a - b))
            elif mode == "ulp":
                v.view(np.uint32)[3 * 4 * 5 + 0] += 1
        if mode == "room_only":               # walls + light: every group is always tested
            keep = np.r_[0:10, 34:36]
            for d_, key in ((osc, "vertices"), (osc, "materials")):
                d_[key] = np.ascontiguousarray(d_[key].reshape(36, 48)[keep]).reshape(-1)
            gsc.vertices = np.ascontiguousarray(gsc.vertices.reshape(36, 48)[keep]).reshape(-1)
            gsc.materials = np.ascontiguousarray(gsc.materials.reshape(36, 48)[keep]).reshape(-1)
            T = 12; osc["T"] = T; gsc.num_triangles = T
        assert np.array_equal(osc["vertices"], gsc.vertices)
        ref, _ = O.pathtrace(osc, W, H, spp)
        r = G.Renderer(gsc, W, H, device=dev)
        for variant in (L.VARIANT_SORTED, L.VARIANT_SORTED_BLOCK):
            img = r.draw(spp=spp, variant=variant)
            assert np.array_equal(img, ref), (mode, variant, int((img != ref).sum()))
        assert np.array_equal(r.draw(spp=spp, variant=L.VARIANT_SMEM), ref), mode


def test_device_object_groups_match_the_cpu_restatement(dev):
    """rt_prep_scene's "group" section (boxes, quad masks, always-tested quads) word for word against
    oracle/rt_groups.c, for the Cornell scene and for scenes with other connectivity."""
    g = np.random.default_rng(5)
    scenes = [O.cornell(64, 64)]
    for mode in ("floating", "clusters", "ulp"):
        sc = O.cornell(64, 64)
        if mode == "ulp":
            sc["vertices"].view(np.uint32)[3 * 4 * 5 + 0] += 1
        else:
            a = None
            for q in range(12):
                if mode == "floating" or q % 4 == 0:
                    a = g.uniform(-2.0, 2.0, 3).astype(np.float32)
                e1 = g.uniform(-0.9, 0.9, 3).astype(np.float32); e2 = g.uniform(-0.9, 0.9, 3).astype(np.float32)
                _quad(sc["vertices"], 10 + 2 * q, a, a + e1, a + e1 + e2, a + e2)
        scenes.append(sc)
    for sc in scenes:
        T, Lc = sc["T"], sc["L"]
        v32 = sc["vertices"].view(np.uint32)
        paired = all(np.array_equal(v32[(3 * i) * 4:(3 * i) * 4 + 3], v32[(3 * (i - 1)) * 4:(3 * (i - 1)) * 4 + 3]) for i in range(1, T, 2))
        inp = np.concatenate([np.array([T, Lc, int(paired), 0], np.uint32).view(np.uint8), sc["vertices"], sc["lights"]])
        got = debug_eval(dev, 6, inp, np.zeros(52, np.uint32))
        want = O.groups(sc)
        assert np.array_equal(got, want), (got, want)


def test_bvh_kernel_small_and_large_scenes(dev):
    """pathTrace_bvh (csrc/bvh.cuh, host_bvh.cpp): every ray query through a bounding-volume
    hierarchy instead of the loop over all triangles.  Same image as the oracle's brute-force loop, bit for bit:
    on the Cornell scene by request (shared edges: the tie rule), and as the library's own choice on a
    tessellation of it with more triangles (1296, 2304) than the brute-force kernels take."""
    from bvh_util import tessellate
    W, H = 96, 64
    for (pos, direction, up, fov), bounces, spp in (((None, None, None, None), 3, 9),
                                                    (((0.3, -0.4, 1.0), (0.2, -0.1, -1.0), (0.0, 1.0, 0.0), 1.6), 4, 5)):
        osc = O.cornell(W, H); gsc = G.init_cornell_box(W, H)
        if pos is not None:
            _set_camera(osc["camera"], pos, direction, up, fov); _set_camera(gsc.camera, pos, direction, up, fov)
        ref, _ = O.pathtrace(osc, W, H, spp, bounces=bounces)
        img = G.Renderer(gsc, W, H, device=dev).draw(spp=spp, bounces=bounces, variant=L.VARIANT_BVH)
        assert np.array_equal(img, ref), (pos, int((img != ref).sum()))
    for n, W, H, spp in ((6, 101, 67, 6), (8, 64, 48, 4)):
        osc = O.cornell(W, H); gsc = G.init_cornell_box(W, H)
        tv, tm, T = tessellate(osc["vertices"], osc["materials"], 36, n)
        osc["vertices"], osc["materials"], osc["T"] = tv, tm, T
        gsc.vertices, gsc.materials, gsc.num_triangles = tv.copy(), tm.copy(), T
        ref, _ = O.pathtrace(osc, W, H, spp)
        r = G.Renderer(gsc, W, H, device=dev)
        img = r.draw(spp=spp)                                           # T > 1024: the library picks the BVH kernel
        assert np.array_equal(img, ref), (n, int((img != ref).sum()))
        assert r.last_command.launch_count() == 3                       # prep, gather, trace
        # progressive: two launches onto one image, same hierarchy reused
        r.clear()
        r.draw_async(spp=spp - 2, firstSample=0, accumulate=1)
        r.draw_async(spp=2, firstSample=spp - 2, accumulate=1).wait_until_completed()
        acc = r.image()
        assert np.array_equal(acc[..., :3] / np.float32(spp), ref[..., :3])


def test_sorted_wide_kernel_few_samples_per_launch(dev):
    """Launches with fewer samples per pixel than a round has slots (progressive rendering) on a groupable scene:
    the library's choice is pathTrace_sortedblk_wide6 (several tiles side by side per warp, rays sorted by object
    group).  Bit-equal to the oracle for 1, 2, 3, 5 spp on a ragged size, as a progressive sum of 1-spp frames,
    and on a scene whose pairing is broken (the kernel's own fall-back to the compacting wide form)."""
    W, H = 203, 75
    osc = O.cornell(W, H)
    r = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev)
    for spp in (1, 2, 3, 5):
        ref, _ = O.pathtrace(osc, W, H, spp)
        img = r.draw(spp=spp)
        assert np.array_equal(img, ref), (spp, int((img != ref).sum()))
    frames = 7
    ref, _ = O.pathtrace(osc, W, H, frames)
    r.clear()
    for k in range(frames):
        cmd = r.draw_async(spp=1, firstSample=k, accumulate=1)
    cmd.wait_until_completed()
    acc = r.image()
    assert np.all(acc[..., 3] == frames) and np.array_equal(acc[..., :3] / np.float32(frames), ref[..., :3])
    osc2 = O.cornell(W, H); gsc2 = G.init_cornell_box(W, H)
    for v in (osc2["vertices"], gsc2.vertices):
        v.view(np.uint32)[3 * 4 * 5 + 0] += 1
    ref2, _ = O.pathtrace(osc2, W, H, 2)
    assert np.array_equal(G.Renderer(gsc2, W, H, device=dev).draw(spp=2), ref2)


def test_prepared_scene_is_reused_until_an_input_changes(dev):
    """rt_prep_scene runs once per scene, not once per command: a second draw is one launch; changing a material
    through contents() + did_modify() runs it again and the image follows (same as a fresh renderer's)."""
    W, H, spp = 96, 64, 8
    osc = O.cornell(W, H)
    ref, _ = O.pathtrace(osc, W, H, spp)
    r = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev)
    assert np.array_equal(r.draw(spp=spp), ref) and r.last_command.launch_count() == 2
    assert np.array_equal(r.draw(spp=spp), ref) and r.last_command.launch_count() == 1
    assert np.array_equal(r.draw(spp=spp, variant=L.VARIANT_SORTED_BLOCK_TILES), ref) and r.last_command.launch_count() == 1
    # the red wall (triangles 2, 3) turns blue
    m = r.materials_buffer.contents(np.float32)
    for t in (2, 3):
        m[t * 12 + 0], m[t * 12 + 2] = 0.0, 0.9
    r.materials_buffer.did_modify()
    om = osc["materials"].view(np.float32)
    for t in (2, 3):
        om[t * 12 + 0], om[t * 12 + 2] = 0.0, 0.9
    ref2, _ = O.pathtrace(osc, W, H, spp)
    img2 = r.draw(spp=spp)
    assert r.last_command.launch_count() == 2
    assert np.array_equal(img2, ref2) and not np.array_equal(ref2, ref)


def test_light_extent_flag_every_kernel_form(dev):
    """RT_VARIANT_LIGHT_EXTENT: sampleAreaLight over the light's own width x depth instead of the
    0.5 x 0.5 of sampling.metal:207-209.  The half extents live in the prepared light (rt_prep_scene), so every kernel
    form takes them from there: each equals the oracle run with the same flag, bit for bit, and -- same renderer, same
    buffers -- goes back to the reference's literals when the flag is dropped (the prepared scene is keyed by it)."""
    W, H, spp = 101, 67, 9
    osc = O.cornell(W, H)
    ref0, _ = O.pathtrace(osc, W, H, spp)
    ref1, cnt1 = O.pathtrace(osc, W, H, spp, light_extent=True)
    assert not np.array_equal(ref0, ref1)
    X = L.VARIANT_LIGHT_EXTENT
    r = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev)
    for variant in (L.VARIANT_DEFAULT, L.VARIANT_SMEM, L.VARIANT_CONST, L.VARIANT_WAVE, L.VARIANT_TILES, L.VARIANT_SORTED,
                    L.VARIANT_SORTED_TILES, L.VARIANT_SORTED_BLOCK, L.VARIANT_SORTED_BLOCK_TILES, L.VARIANT_BVH):
        assert np.array_equal(r.draw(spp=spp, variant=variant | X), ref1), variant
        assert np.array_equal(r.draw(spp=spp, variant=variant), ref0), variant
    assert np.array_equal(r.draw(spp=spp, variant=X), ref1) and r.last_command.launch_count() == 2     # flag changed: prep again
    assert np.array_equal(r.draw(spp=spp, variant=X), ref1) and r.last_command.launch_count() == 1
    c = r.count_rays(spp=spp, variant=X)
    assert [c["paths"], c["closest"], c["shadow"], c["shaded"]] == [int(v) for v in cnt1[:4]]
    for few in (1, 3):                                                 # the wide forms (fewer samples than a round has slots)
        refw, _ = O.pathtrace(osc, W, H, few, light_extent=True)
        assert np.array_equal(r.draw(spp=few, variant=X), refw), few
        assert np.array_equal(r.draw(spp=few, variant=L.VARIANT_WIDE | X), refw), few
    # config 4's scene with four lights of four sizes (one of them the literal 0.5 x 0.5)
    W2, H2 = 120, 68
    osc = O.spheres(W2, H2); gsc = G.init_sphere_scene(W2, H2)
    for l, wd in enumerate(((0.9, 0.3), (0.2, 1.1), (0.5, 0.5), (1.4, 1.4))):
        osc["lights"].view(np.float32)[16 * l + 12:16 * l + 14] = wd
        gsc.lights.view(np.float32)[16 * l + 12:16 * l + 14] = wd
    ref, _ = O.pathtrace(osc, W2, H2, 4, light_extent=True)
    lit, _ = O.pathtrace(osc, W2, H2, 4)
    assert not np.array_equal(ref, lit)
    r2 = G.Renderer(gsc, W2, H2, device=dev)
    for variant in (L.VARIANT_DEFAULT, L.VARIANT_SMEM, L.VARIANT_CONST, L.VARIANT_WAVE):
        assert np.array_equal(r2.draw(spp=4, variant=variant | X), ref), variant
    assert np.array_equal(r2.draw(spp=4), lit)
    # a tessellated Cornell scene above the brute-force limit: the BVH kernel by the library's own choice
    from bvh_util import tessellate
    W3, H3 = 64, 48
    osc = O.cornell(W3, H3); gsc = G.init_cornell_box(W3, H3)
    tv, tm, T = tessellate(osc["vertices"], osc["materials"], 36, 6)
    osc["vertices"], osc["materials"], osc["T"] = tv, tm, T
    gsc.vertices, gsc.materials, gsc.num_triangles = tv.copy(), tm.copy(), T
    ref, _ = O.pathtrace(osc, W3, H3, 4, light_extent=True)
    assert np.array_equal(G.Renderer(gsc, W3, H3, device=dev).draw(spp=4, variant=X), ref)


@pytest.mark.parametrize("mode", ["stretched", "flush"])
def test_exit_and_light_plane_rules_in_other_rooms(dev, mode):
    """The block-sorted kernel's exit rule and light-plane rule away from the Cornell room's round numbers: the scene under an
    affine map with three factors, and the light raised into the ceiling's plane (tests/test_group_cull_cpu.py: other_room, the
    CPU checker's side of the same scenes); a camera outside and one inside the room, every sorted form and the MIS twin."""
    from test_group_cull_cpu import other_room
    W, H, spp = 176, 120, 12
    for cam in (None, (0.3, -0.4, 1.5)):
        osc = other_room(W, H, mode)
        if cam is not None:
            c = osc["camera"].view(np.float32)
            c[0:3] = cam; c[4:7] = (0.1, -0.1, -1.0); c[14] = 1.4
        gsc = G.init_cornell_box(W, H)
        gsc.vertices[:] = osc["vertices"].view(np.uint8); gsc.lights[:] = osc["lights"].view(np.uint8); gsc.camera[:] = osc["camera"].view(np.uint8)
        ref, _ = O.pathtrace(osc, W, H, spp, bounces=4)
        r = G.Renderer(gsc, W, H, device=dev)
        base = r.draw(spp=spp, bounces=4, variant=L.VARIANT_SMEM)
        assert np.array_equal(base, ref), (mode, cam)
        for variant in (L.VARIANT_DEFAULT, L.VARIANT_SORTED, L.VARIANT_SORTED_BLOCK, L.VARIANT_SORTED_BLOCK_TILES):
            img = r.draw(spp=spp, bounces=4, variant=variant)
            assert np.array_equal(img, ref), (mode, cam, variant, int((img != ref).sum()))
        refm, _ = O.pathtrace(osc, W, H, spp, bounces=4, mis=O.MIS)
        assert np.array_equal(r.draw(spp=spp, bounces=4, variant=L.VARIANT_SORTED_BLOCK | L.VARIANT_MIS), refm), (mode, cam)
        assert np.array_equal(r.draw(spp=3, bounces=4), O.pathtrace(osc, W, H, 3, bounces=4)[0]), (mode, cam)      # the wide form
