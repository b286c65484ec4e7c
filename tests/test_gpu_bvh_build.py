"""The acceleration structure built ON THE DEVICE (csrc/bvh_build.cuh; reference: setupAccelerationStructures,
RTrace/computeShader.swift:45-97, which builds on the GPU at init): Morton keys, bitonic sort, Karras radix tree, bottom-up
boxes, leaves of at most 12.  Checked: the trees are valid (the structural checker of the host builder's CPU test), the
build takes under 2 ms at 36 864 triangles, and images traced through them -- via the handle made by
rtAccelerationStructureCreate and via the queue's own lazy build -- are the brute-force oracle's bit for bit."""
import ctypes as C

import numpy as np
import pytest

import gpuraytracer_b200 as G
from gpuraytracer_b200 import _lib as L
import oracle_lib as O
from bvh_util import tessellate, check_nodes

pytestmark = pytest.mark.gpu
L.lib.rtDebugBuildBvhDevice.restype = C.c_int
L.lib.rtDebugBuildBvhDevice.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p,
                                        C.POINTER(C.c_uint32), C.POINTER(C.c_float), C.POINTER(C.c_uint32)]


@pytest.fixture(scope="module")
def dev():
    return G.Device(0)


def device_build(dev, verts, T):
    nodes = np.zeros((max(T, 2), 16), np.uint32)
    order = np.zeros(T, np.uint32)
    n, ms, st = C.c_uint32(), C.c_float(), C.c_uint32()
    v = np.ascontiguousarray(verts)
    L.check(L.lib.rtDebugBuildBvhDevice(dev.h, v.ctypes.data_as(C.c_void_p), T, nodes.ctypes.data_as(C.c_void_p), nodes.shape[0],
                                        order.ctypes.data_as(C.c_void_p), C.byref(n), C.byref(ms), C.byref(st)), "rtDebugBuildBvhDevice")
    return nodes[:n.value], order, ms.value, st.value


def test_device_built_trees_are_valid(dev):
    sc = O.cornell(64, 64)
    for T in (1, 2, 3, 5, 12, 13, 25, 36):
        nodes, order, ms, st = device_build(dev, sc["vertices"][:48 * T].copy(), T)
        check_nodes(nodes, order, sc["vertices"][:48 * T].copy(), T)
        assert st == 0
    same = np.tile(sc["vertices"][:48], 37).copy()                 # 37 copies of one triangle: equal Morton codes, ids break the tie
    nodes, order, ms, st = device_build(dev, same, 37)
    check_nodes(nodes, order, same, 37)
    report = []
    for n in (6, 8, 32):                                           # 1 296, 2 304, 36 864 triangles
        tv, tm, T = tessellate(sc["vertices"], sc["materials"], 36, n)
        nodes, order, ms, st = device_build(dev, tv, T)
        depth, leaves, per_leaf = check_nodes(nodes, order, tv, T)
        assert st == 0 and depth <= 40 and per_leaf > 3.0, (T, depth, per_leaf)
        report.append((T, ms, depth, leaves, per_leaf))
    print("device BVH builds (T, ms, depth, leaves, triangles per leaf):", report)
    assert report[-1][1] < 2.0, report                             # build < 2 ms at 36 864 triangles


def _tess_scene(W, H, n):
    sc, osc = G.init_cornell_box(W, H), O.cornell(W, H)
    tv, tm, T = tessellate(sc.vertices, sc.materials, 36, n)
    sc.vertices, sc.materials, sc.num_triangles = tv, tm, T
    return sc, dict(osc, vertices=tv, materials=tm, T=T)


@pytest.mark.parametrize("n", [6, 8])
def test_images_through_device_built_trees(dev, n):
    """Renderer makes the handle (rtAccelerationStructureCreate builds the tree at once, as the reference does at init) and
    binds it; the image is the oracle's brute force over the same triangles."""
    W, H, spp = 80, 60, 4
    sc, osc = _tess_scene(W, H, n)
    ref, _ = O.pathtrace(osc, W, H, spp)
    r = G.Renderer(sc, W, H, device=dev)
    img = r.draw(spp=spp)
    assert np.array_equal(img, ref), int((img != ref).sum())
    assert r.last_command.launch_count() <= 3                     # prep + gather + trace: no build inside the command
    # the vertices change: the handle's tree is stale, the queue rebuilds its own on the device (6 more launches), same bits
    v = r.vertex_buffer.contents(np.float32)
    v[:] = v                                                        # same values, new epoch
    r.vertex_buffer.did_modify()
    img2 = r.draw(spp=spp)
    assert np.array_equal(img2, ref)
    assert r.last_command.launch_count() >= 8
    assert np.array_equal(r.draw(spp=spp), ref) and r.last_command.launch_count() <= 2      # and keeps it


def test_36864_triangles_rows_against_brute_force(dev):
    W, H, spp = 160, 120, 2
    sc, osc = _tess_scene(W, H, 32)
    r = G.Renderer(sc, W, H, device=dev)
    img = r.draw(spp=spp)
    for row0 in (10, 60, 100):
        ref, _ = O.pathtrace(osc, W, H, spp, row0=row0, rows=2)
        assert np.array_equal(img[row0:row0 + 2], ref), row0


def test_a_million_triangles(dev):
    """The advertised range (RT_MAX_TRIANGLES_BVH = 4 M): the Cornell scene cut into 36 x 167 x 167 = 1 004 004 triangles.
    rt_prep_scene runs as a grid, the tree is built on the device, the image of sampled rows is the oracle's brute force."""
    from bvh_util import tessellate_fast
    W, H, spp = 64, 48, 1
    sc, osc = G.init_cornell_box(W, H), O.cornell(W, H)
    tv, tm, T = tessellate_fast(sc.vertices, sc.materials, 36, 167)
    assert T == 1004004
    sc.vertices, sc.materials, sc.num_triangles = tv, tm, T
    osc = dict(osc, vertices=tv, materials=tm, T=T)
    nodes, order, ms, st = device_build(dev, tv, T)
    assert st == 0 and ms < 20.0, (ms, st)
    print(f"device BVH build at {T} triangles: {ms:.2f} ms")
    r = G.Renderer(sc, W, H, device=dev)
    img = r.draw(spp=spp)
    ms_draw = r.last_command.gpu_time_ms()
    for row0 in (12, 30):
        ref, _ = O.pathtrace(osc, W, H, spp, row0=row0, rows=1)
        assert np.array_equal(img[row0:row0 + 1], ref), row0
    print(f"  {W}x{H}x{spp} through it: {ms_draw:.2f} ms")
