"""``draw_triangle``: the reference's mode-B host driver
(Sources/gpuRaytracer/computeShader.swift:99-189, called from main.swift:84) over the C ABI:
pipeline "drawTriangle", buffers 0-4 and 6 at the reference's indices, acceleration structure
at 5, one thread per pixel, commit, wait, read the RGBA8 pixels and the float3 debug buffer back.
Nothing is computed in Python."""
import ctypes as C

import numpy as np

from . import _lib as L
from .api import Command, Device, default_params
from .scene import SceneBytes, init_cornell_box
from ._lib import check, lib


def draw_triangle(scene: SceneBytes = None, width=800, height=600, *, device=None, ordinal=0, row0=0, rows=None,
                  camera_rays=6, mis_samples=300, tile_rows=0, tile_stride=0, return_command=False):
    """Returns (pixels uint8 [rows,width,4], text float32 [rows,width,3]) for global rows [row0, row0+rows)."""
    dev = device or Device(ordinal)
    scene = scene or init_cornell_box(width, height, light_size=1.5)      # main.swift:23-66: 1.5 x 1.5 light
    rows = height - row0 if rows is None else rows
    npx = width * rows
    pipeline = dev.make_pipeline("drawTriangle")                          # computeShader.swift:112-123
    queue = dev.make_queue()                                              # :125
    camera_buffer = dev.make_buffer(scene.camera)                         # :138-149
    materials_buffer = dev.make_buffer(scene.materials)
    pixels_buffer = dev.make_buffer(length=npx * 4)
    lights_buffer = dev.make_buffer(scene.lights)
    vertex_buffer = dev.make_buffer(scene.vertices)
    debug_buffer = dev.make_buffer(length=npx * 16)                       # :148 over-allocates 16 B per pixel
    accel = C.c_void_p()
    check(lib.rtAccelerationStructureCreate(dev.h, vertex_buffer.h, 0, 16, scene.num_triangles, C.byref(accel)),
          "rtAccelerationStructureCreate")
    try:
        cmd = Command(queue, pipeline)
        cmd.set_buffer(camera_buffer, 0, L.INDEX_CAMERA)                  # :153-159
        cmd.set_buffer(materials_buffer, 0, L.INDEX_MATERIALS)
        cmd.set_buffer(lights_buffer, 0, L.INDEX_LIGHTS)
        cmd.set_buffer(vertex_buffer, 0, L.INDEX_VERTICES)
        cmd.set_buffer(pixels_buffer, 0, L.INDEX_PIXELS)
        check(lib.rtSetAccelerationStructure(cmd.h, accel, L.INDEX_ACCEL), "rtSetAccelerationStructure")
        cmd.set_buffer(debug_buffer, 0, L.INDEX_DEBUG)
        cmd.set_params(default_params(width, height, row0=row0, rows=rows, numTriangles=scene.num_triangles,
                                      numLights=scene.num_lights, cameraRays=camera_rays, misSamples=mis_samples,
                                      tileRows=tile_rows, tileStride=tile_stride))
        cmd.dispatch_threads((width, rows, 1), (8, 8, 1))                 # :164-170
        cmd.commit()
        cmd.wait_until_completed()                                        # :172-174
        text = debug_buffer.contents(np.float32)[:npx * 3].reshape(rows, width, 3).copy()     # :176-182
        pixels = pixels_buffer.contents(np.uint8).reshape(rows, width, 4).copy()              # :184-188
    finally:
        lib.rtAccelerationStructureRelease(accel)
    return (pixels, text, cmd) if return_command else (pixels, text)
