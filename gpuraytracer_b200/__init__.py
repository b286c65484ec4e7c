"""gpuraytracer_b200 -- the per-pixel path-tracing kernel of Nishad-Sharma/gpuRaytracer
as hand-written CUDA for sm_100a behind a C ABI (include/rtcuda.h).

    csrc/pathtrace.cu     the kernels
    csrc/rtcuda.cpp       the C ABI over the CUDA driver API  -> librtcuda.so
    csrc/host_scene.cpp   C++ mirror of scene.swift / computeShader.swift converters
    api.py, renderer.py   ctypes objects and the Renderer mirror of renderer.swift
    compute_shader.py     draw_triangle: mirror of the mode-B host driver (computeShader.swift:99-189)
    distributed.py        row tiles across ranks + NCCL gather
"""
from ._lib import (RTError, RTLaunchParams, VARIANT_CONST, VARIANT_DEFAULT, VARIANT_SMEM, VARIANT_WAVE)  # noqa: F401
from .api import Buffer, Command, Device, Pipeline, Queue, default_params  # noqa: F401
from .renderer import Renderer  # noqa: F401
from .compute_shader import draw_triangle  # noqa: F401
from .scene import SceneBytes, fill_offsets, init_cornell_box, init_sphere_scene  # noqa: F401
