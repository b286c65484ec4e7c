"""rtPathTraceNoAlloc (include/rtcuda.h; csrc/noalloc.cu): pathTrace for callers that cannot allocate device memory.  Its
scratch (the prepared scene) is a static __device__ array inside the library, reached through cudaGetSymbolAddress.
No counterpart in the reference.

Checked here, through the C ABI, against the existing entry point (a pathTrace command, rtCommit) on the same inputs:
the images must be bit-equal -- the tolerance every parity test of this repo uses -- on two calls in a row (the second
call finds the first call's contents in the shared scratch), for a second scene of another size after that (other
layout in the same array), and the limits must come back as error codes.  This file is collected last on purpose:
a failure here must not hide the other tests.  It passes alone in a fresh process and
after the others."""
import ctypes as C

import numpy as np
import pytest

import gpuraytracer_b200 as G
from gpuraytracer_b200 import _lib as L
from gpuraytracer_b200.api import Buffer
from gpuraytracer_b200._lib import lib

pytestmark = pytest.mark.gpu


class Inputs:
    """Device buffers made by the existing API (rtBufferCreate); the new entry point gets their device pointers."""

    def __init__(self, dev, scene, W, H, seed=1):
        self.scene, self.W, self.H = scene, W, H
        self.offsets = G.fill_offsets(W, 0, H, seed=seed)
        self.bufs = [dev.make_buffer(x) for x in (scene.camera, scene.materials, scene.lights, scene.vertices)]
        self.spheres = dev.make_buffer(scene.spheres) if scene.num_spheres else None
        self.off = dev.make_buffer(self.offsets)
        self.out = dev.make_buffer(length=W * H * 16)
        # a second handle on the output's memory that the library does not own: its contents() is a device read every time
        self.view = Buffer(dev, length=W * H * 16, device_pointer=self.out.device_pointer)

    def params(self, **over):
        s = self.scene
        return G.default_params(self.W, self.H, row0=0, rows=self.H, numTriangles=s.num_triangles,
                                numSpheres=s.num_spheres, numLights=s.num_lights, **over)

    def no_alloc(self, **over):
        p = self.params(**over)
        rc = lib.rtPathTraceNoAlloc(*(C.c_void_p(b.device_pointer) for b in self.bufs),
                                    C.c_void_p(self.spheres.device_pointer if self.spheres else None),
                                    C.c_void_p(self.off.device_pointer), C.c_void_p(self.out.device_pointer), C.byref(p), None)
        return rc

    def image(self):
        # launched on the NULL stream; the device-to-host copy behind contents() is ordered after it
        return self.view.contents(np.float32).reshape(self.H, self.W, 4).copy()

    def poison(self):
        self.out.contents(np.float32)[:] = np.float32(-7.0)
        self.out.did_modify()


def existing_entry_point(dev, scene, W, H, **over):
    return G.Renderer(scene, W, H, device=dev).draw(variant=L.VARIANT_WAVE, **over)


def test_no_alloc_matches_the_command_path_two_calls_in_a_row_then_another_scene():
    dev = G.Device(0)                                      # makes the device's primary context current on this thread
    W, H, spp = 200, 77, 13                                # ragged against the 16 x 8 block; an odd sample count
    cornell = Inputs(dev, G.init_cornell_box(W, H), W, H)
    want = existing_entry_point(dev, cornell.scene, W, H, spp=spp)
    for call in (1, 2):
        cornell.poison()
        assert cornell.no_alloc(spp=spp) == L.RT_OK, lib.rtLastErrorString()
        got = cornell.image()
        assert np.array_equal(got, want), (call, int((got != want).sum()))
    # another layout in the same scratch: 18 triangles + 256 spheres + 4 lights (the R = 4 form), light-extent flag on
    W2, H2 = 96, 64
    sph = Inputs(dev, G.init_sphere_scene(W2, H2), W2, H2)
    want2 = G.Renderer(sph.scene, W2, H2, device=dev).draw(spp=5, variant=L.VARIANT_WAVE | L.VARIANT_LIGHT_EXTENT)
    assert sph.no_alloc(spp=5, variant=L.VARIANT_LIGHT_EXTENT) == L.RT_OK, lib.rtLastErrorString()
    assert np.array_equal(sph.image(), want2)
    # and back: the Cornell scene finds the sphere scene's words in the scratch
    cornell.poison()
    assert cornell.no_alloc(spp=spp, bounces=2) == L.RT_OK
    assert np.array_equal(cornell.image(), existing_entry_point(dev, cornell.scene, W, H, spp=spp, bounces=2))


def test_no_alloc_limits_are_error_codes():
    dev = G.Device(0)
    W, H = 64, 48
    inp = Inputs(dev, G.init_cornell_box(W, H), W, H)
    inp.poison()
    p = inp.params(spp=4)
    p.numTriangles = 257
    assert lib.rtPathTraceNoAlloc(*(C.c_void_p(b.device_pointer) for b in inp.bufs), None, C.c_void_p(inp.off.device_pointer),
                                  C.c_void_p(inp.out.device_pointer), C.byref(p), None) == -5
    assert b"at most 256 triangles" in lib.rtLastErrorString()
    for over, code in ((dict(numTriangles=257), -5), (dict(numSpheres=257), -5), (dict(numLights=17), -5),
                       (dict(bounces=5), -7), (dict(variant=L.VARIANT_SMEM), -7), (dict(spp=0), -1)):
        p = inp.params(spp=4)
        for k, v in over.items():
            setattr(p, k, v)
        rc = lib.rtPathTraceNoAlloc(*(C.c_void_p(b.device_pointer) for b in inp.bufs), C.c_void_p(inp.off.device_pointer),
                                    C.c_void_p(inp.off.device_pointer), C.c_void_p(inp.out.device_pointer), C.byref(p), None)
        assert rc == code, (over, rc, lib.rtLastErrorString())
    p = inp.params(spp=4)
    assert lib.rtPathTraceNoAlloc(None, None, None, None, None, None, None, C.byref(p), None) == -1
    assert np.all(inp.image() == np.float32(-7.0))         # a rejected call launches nothing
