"""``Renderer``: the reference's host renderer (RTrace/renderer.swift:9-146) over the
C ABI.  ``Renderer(...)`` does what ``Renderer.init`` does (device, pipeline, queue,
scene buffers, the two "textures", the acceleration-structure record); ``draw()``
does what ``Renderer.draw`` does (bind, dispatch, commit, wait) and returns the
float4 image instead of writing a PNG (``save_png`` covers image.swift:15-100).

The scene, resolution and the shader's literals (400 spp, 3 bounces) default to the
reference's; BASELINE.json's configs change them through keyword arguments.
"""
import ctypes as C

import numpy as np

from . import _lib as L
from .api import Buffer, Command, Device, default_params, _vp
from .scene import SceneBytes, fill_offsets, init_cornell_box
from ._lib import check, lib


class Renderer:
    def __init__(self, scene: SceneBytes = None, width=800, height=600, *, device=None, ordinal=0,
                 kernel="pathTrace", row0=0, rows=None, seed=1, offsets=None, stream=None,
                 output_device_pointer=None):
        self.device = device or Device(ordinal)                       # renderer.swift:30
        self.pipeline = self.device.make_pipeline(kernel)             # :32-43
        self.queue = self.device.make_queue(stream)                   # :44
        self.scene = scene or init_cornell_box(width, height)         # :46
        self.width, self.height = width, height
        self.row0 = row0
        self.rows = height - row0 if rows is None else rows
        s = self.scene
        self.camera_buffer = self.device.make_buffer(s.camera)        # :61-72
        self.materials_buffer = self.device.make_buffer(s.materials)
        self.lights_buffer = self.device.make_buffer(s.lights)
        self.vertex_buffer = self.device.make_buffer(s.vertices)
        self.spheres_buffer = self.device.make_buffer(s.spheres) if s.num_spheres else None
        npx = self.width * self.rows
        # renderTexture / randomTexture (:74-110) as linear buffers holding this tile's rows
        if output_device_pointer is not None:
            self.render_buffer = Buffer(self.device, length=npx * 16, device_pointer=output_device_pointer)
        else:
            self.render_buffer = self.device.make_buffer(length=npx * 16)
        if offsets is None:
            offsets = fill_offsets(width, row0, self.rows, seed)
        assert offsets.dtype == np.uint32 and offsets.size == npx
        self.random_buffer = self.device.make_buffer(offsets)
        accel = C.c_void_p()                                          # :113, computeShader.swift:45-97
        check(lib.rtAccelerationStructureCreate(self.device.h, self.vertex_buffer.h, 0, 16, s.num_triangles,
                                                C.byref(accel)), "rtAccelerationStructureCreate")
        self.accel = accel
        self.last_command = None
        self.debug_buffer = None
        self.light_normals_buffer = None

    def set_light_normals(self, normals):
        """Bind a per-light normal table (float32 [L, 4], xyz = normal) at RT_INDEX_LIGHT_NORMALS, or None to go back to
        the reference's literal (0,-1,0) (sampling.metal:205).  SURVEY N3."""
        if normals is None:
            self.light_normals_buffer = None
            return
        n = np.ascontiguousarray(normals, np.float32).reshape(-1, 4)
        assert n.shape[0] == self.scene.num_lights
        self.light_normals_buffer = self.device.make_buffer(n)

    def params(self, **over):
        p = default_params(self.width, self.height, row0=self.row0, rows=self.rows,
                           numTriangles=self.scene.num_triangles, numSpheres=self.scene.num_spheres,
                           numLights=self.scene.num_lights)
        for k, v in over.items():
            setattr(p, k, v)
        return p

    def encode(self, **over):
        """renderer.swift:118-141: bind everything and dispatch; returns the un-committed command."""
        cmd = Command(self.queue, self.pipeline)
        cmd.set_buffer(self.camera_buffer, 0, L.INDEX_CAMERA)
        cmd.set_buffer(self.materials_buffer, 0, L.INDEX_MATERIALS)
        cmd.set_buffer(self.lights_buffer, 0, L.INDEX_LIGHTS)
        cmd.set_buffer(self.vertex_buffer, 0, L.INDEX_VERTICES)
        check(lib.rtSetAccelerationStructure(cmd.h, self.accel, L.INDEX_ACCEL), "rtSetAccelerationStructure")
        if self.spheres_buffer is not None:
            cmd.set_buffer(self.spheres_buffer, 0, L.INDEX_SPHERES)
        if self.light_normals_buffer is not None:
            cmd.set_buffer(self.light_normals_buffer, 0, L.INDEX_LIGHT_NORMALS)
        cmd.set_buffer(self.render_buffer, 0, L.INDEX_OUTPUT)          # setTexture(renderTexture, 0)
        cmd.set_buffer(self.random_buffer, 0, L.INDEX_OFFSETS)         # setTexture(randomTexture, 1)
        params = self.params(**over)
        if params.variant & L.VARIANT_COUNT_RAYS:
            # the reference binds a debug buffer at index 6 and never writes it (renderer.swift:128);
            # the counting variant uses it for the exact ray counters
            if self.debug_buffer is None:
                self.debug_buffer = self.device.make_buffer(length=32)
            cmd.set_buffer(self.debug_buffer, 0, L.INDEX_DEBUG)
        cmd.set_params(params)
        cmd.dispatch_threads((self.width, self.rows, 1), (8, 8, 1))    # :137-141
        return cmd

    def draw_async(self, **over):
        cmd = self.encode(**over)
        cmd.commit()                                                   # :143
        self.last_command = cmd
        return cmd

    def draw(self, **over):
        """Render and return the tile as float32 [rows, width, 4] (a copy)."""
        cmd = self.draw_async(**over)
        cmd.wait_until_completed()                                     # :144
        return self.image()

    def image(self):
        return self.render_buffer.contents(np.float32).reshape(self.rows, self.width, 4).copy()

    def count_rays(self, **over):
        """Render once with the counting variant; returns {paths, closest, shadow, shaded} (exact)."""
        if self.debug_buffer is not None:
            self.debug_buffer.contents(np.uint64)[:] = 0
            self.debug_buffer.did_modify()
        v, flags = over.get("variant", 0) & 0xff, over.get("variant", 0) & L.VARIANT_LIGHT_EXTENT
        over["variant"] = (v if v <= L.VARIANT_WAVE else 0) | L.VARIANT_COUNT_RAYS | flags   # tuning variants count with the default kernel
        self.draw_async(**over).wait_until_completed()
        c = self.debug_buffer.contents(np.uint64)
        return dict(paths=int(c[0]), closest=int(c[1]), shadow=int(c[2]), shaded=int(c[3]))

    def clear(self):
        """Zero the accumulation target (progressive mode starts from 0)."""
        v = self.render_buffer.contents(np.float32)
        v[:] = 0
        self.render_buffer.did_modify()

    def resolve_rgba8(self, divide_by_alpha=False):
        """image.swift:40-65 on the device; returns uint8 [rows, width, 4]."""
        npx = self.width * self.rows
        out = self.device.make_buffer(length=npx * 4)
        check(lib.rtResolveToRGBA8(self.queue.h, self.render_buffer.h, 0, out.h, 0, npx, int(divide_by_alpha)),
              "rtResolveToRGBA8")
        return out.contents(np.uint8).reshape(self.rows, self.width, 4).copy()

    def resolve_rgba16f(self, divide_by_alpha=False):
        """The image in the reference's render-target format (rgba16Float, renderer.swift:74-82); float16 [rows, width, 4]."""
        npx = self.width * self.rows
        out = self.device.make_buffer(length=npx * 8)
        check(lib.rtResolveToRGBA16F(self.queue.h, self.render_buffer.h, 0, out.h, 0, npx, int(divide_by_alpha)),
              "rtResolveToRGBA16F")
        return out.contents(np.float16).reshape(self.rows, self.width, 4).copy()

    def save_png(self, path, divide_by_alpha=False):
        from PIL import Image
        Image.fromarray(self.resolve_rgba8(divide_by_alpha), "RGBA").save(path)

    def release(self):
        if self.accel:
            lib.rtAccelerationStructureRelease(self.accel)
            self.accel = None
