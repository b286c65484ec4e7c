"""Which kernel the library picks (csrc/rt_plan.cpp's rule table) is pinned here for every BASELINE.json config on one
B200 and for one rank's share of an 8-GPU frame, so that a threshold edit cannot silently change the headline kernel.
The choice is a pure function of the launch parameters and three device numbers: no GPU involved.  (The reference has one
kernel, named in makeFunction(name:), RTrace/renderer.swift:39; the forms are schedules of it with bit-identical output.)"""
import pytest

import gpuraytracer_b200 as G
from gpuraytracer_b200 import _lib as L
from gpuraytracer_b200 import distributed as D
from gpuraytracer_b200.api import choose_kernel, default_params


def cornell(width, height, **over):
    return default_params(width, height, row0=0, rows=height, numTriangles=36, numSpheres=0, numLights=1, **over)


def test_config1_cornell_512_16spp():
    c = choose_kernel(cornell(512, 512, spp=16))
    assert c["kernel"] == "pathTrace_sortedblk6" and c["grid"] == (32, 64)


def test_config2_headline_1080p_1024spp():
    c = choose_kernel(cornell(1920, 1080, spp=1024))
    assert c["kernel"] == "pathTrace_sortedblk6" and c["grid"] == (120, 135)
    # scene + 768 slots of 15.5 words + lists: four blocks per SM
    assert 4 * (c["smem_bytes"] + 1024) <= 233472 - 1024 < 5 * (c["smem_bytes"] + 1024)


def test_config3_4k_4096spp_one_gpu_and_shares():
    assert choose_kernel(cornell(3840, 2160, spp=4096))["kernel"] == "pathTrace_sortedblk6"
    for world in (2, 4, 8):
        row0, rows, tr, ts = D.tile_params(2160, 0, world, 8)
        p = default_params(3840, 2160, row0=row0, rows=rows, numTriangles=36, numLights=1, spp=4096, tileRows=tr, tileStride=ts)
        # 2160 / 8 = 270 rows = 8 160 blocks = 13.8 resident waves: the plain grid
        assert choose_kernel(p)["kernel"] == "pathTrace_sortedblk6", world


def test_config2_one_rank_of_8_takes_the_ticketed_form():
    row0, rows, tr, ts = D.tile_params(1080, 0, 8, 8)
    p = default_params(1920, 1080, row0=row0, rows=rows, numTriangles=36, numLights=1, spp=1024, tileRows=tr, tileStride=ts)
    c = choose_kernel(p)
    assert c["kernel"] == "pathTrace_sortedblk_tiles6"
    assert c["chunks"] == 8 and c["chunk_spp"] == 132 and c["grid"] == (120, 17)      # 2 040 tiles x 8 chunks
    for world, kernel in ((2, "pathTrace_sortedblk6"), (4, "pathTrace_sortedblk_tiles6")):
        row0, rows, tr, ts = D.tile_params(1080, 0, world, 8)
        p = default_params(1920, 1080, row0=row0, rows=rows, numTriangles=36, numLights=1, spp=1024, tileRows=tr, tileStride=ts)
        assert choose_kernel(p)["kernel"] == kernel, world


def test_config4_spheres():
    p = default_params(1920, 1080, row0=0, rows=1080, numTriangles=18, numSpheres=256, numLights=4, spp=256)
    c = choose_kernel(p)
    assert c["kernel"] == "pathTrace_wave4"          # the 25 KB scene leaves room for four R = 4 blocks, not four R = 6


def test_config5_progressive_one_sample_per_frame():
    p = cornell(1920, 1080, spp=1, firstSample=17, accumulate=1)
    c = choose_kernel(p)
    assert c["kernel"] == "pathTrace_sortedblk_wide6" and c["grid"] == (20, 135)      # six 16 x 8 tiles side by side per block


def test_mode_b_and_large_scenes():
    assert choose_kernel(cornell(512, 512), kind=1)["kernel"] == "drawTriangle_mis"
    p = default_params(640, 480, row0=0, rows=480, numTriangles=36864, numLights=1, spp=16)
    assert choose_kernel(p)["kernel"] == "pathTrace_bvh"
    p.numTriangles = 1 << 23
    with pytest.raises(G.RTError):
        choose_kernel(p)


def test_requests_by_name_and_their_fallbacks():
    base = dict(spp=64)
    name = lambda **kw: choose_kernel(cornell(1920, 1080, **{**base, **kw}))["kernel"]
    assert name(variant=L.VARIANT_SMEM) == "pathTrace_smem"
    assert name(variant=L.VARIANT_CONST) == "pathTrace_const"
    assert name(variant=L.VARIANT_WAVE) == "pathTrace_wave"
    assert name(variant=L.VARIANT_TILES) == "pathTrace_tiles"
    assert name(variant=L.VARIANT_SORTED) == "pathTrace_sorted6"
    assert name(variant=L.VARIANT_SORTED_TILES) == "pathTrace_sorted_tiles6"
    assert name(variant=L.VARIANT_SORTED_BLOCK) == "pathTrace_sortedblk6"
    assert name(variant=L.VARIANT_SORTED_BLOCK_TILES) == "pathTrace_sortedblk_tiles6"
    assert name(variant=L.VARIANT_BVH) == "pathTrace_bvh"
    assert name(variant=L.VARIANT_COUNT_RAYS) == "pathTrace_wave_count"
    assert name(variant=L.VARIANT_COUNT_RAYS, spp=2) == "pathTrace_smem_count"
    assert name(variant=L.VARIANT_WIDE) == "pathTrace_wave"                  # the wide form is for spp < 6
    assert name(variant=L.VARIANT_WIDE, spp=2) == "pathTrace_wide"
    assert name(variant=L.VARIANT_SORTED_BLOCK, spp=5) == "pathTrace_wave"     # fewer samples than a sorted round
    # an odd triangle count or a second light cannot be grouped: the compacting kernel
    p = cornell(1920, 1080, spp=64); p.numTriangles = 35
    assert choose_kernel(p)["kernel"] == "pathTrace_wave"
    p = cornell(1920, 1080, spp=64); p.numLights = 2
    assert choose_kernel(p)["kernel"] == "pathTrace_wave"
    with pytest.raises(G.RTError):
        name(variant=9)                                                          # pathTrace_exp9 exists in tuning builds only
