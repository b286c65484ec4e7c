"""Times the BVH kernel (csrc/bvh.cuh) on tessellations of the Cornell scene: python tools/time_bvh.py [W H SPP]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import gpuraytracer_b200 as G
from gpuraytracer_b200 import _lib as L
from bvh_util import tessellate

W, H, SPP = (int(a) for a in sys.argv[1:4]) if len(sys.argv) >= 4 else (1920, 1080, 16)
dev = G.Device(0)
base = G.init_cornell_box(W, H)
ref = None
for n in (1, 4, 8, 16, 32):
    sc = G.init_cornell_box(W, H)
    if n > 1:
        tv, tm, T = tessellate(base.vertices, base.materials, 36, n)
        sc.vertices, sc.materials, sc.num_triangles = tv, tm, T
    r = G.Renderer(sc, W, H, device=dev)
    t0 = time.time(); r.draw(spp=2, variant=L.VARIANT_BVH); build_s = time.time() - t0     # includes the host build
    ts = []
    for _ in range(3):
        r.draw_async(spp=SPP, variant=L.VARIANT_BVH).wait_until_completed()
        ts.append(r.last_command.gpu_time_ms())
    ms = min(ts)
    line = f"T = {sc.num_triangles:6d}: {ms:8.2f} ms  {W * H * SPP / ms / 1e3:8.1f} Mpaths/s  (first draw incl. host build {build_s * 1e3:.0f} ms)"
    if n == 1:
        r.draw_async(spp=SPP).wait_until_completed()
        line += f" | brute-force default kernel {r.last_command.gpu_time_ms():.2f} ms"
    print(line, flush=True)
