"""Times the BVH kernel (csrc/bvh.cuh) on tessellations of the Cornell scene: python tools/time_bvh.py [W H SPP]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import gpuraytracer_b200 as G
from gpuraytracer_b200 import _lib as L
from bvh_util import tessellate

W, H, SPP = (int(a) for a in sys.argv[1:4]) if len(sys.argv) >= 4 else (1920, 1080, 16)
# BVH_FORMS="-,l6,s4,...": kernel forms of a tuning build (RT_BVH_FORM, csrc/bvh.cuh), "-" = the shipped kernel; each form's
# image is compared with the shipped kernel's (which the GPU tests compare with the oracle)
FORMS = [f for f in os.environ.get("BVH_FORMS", "-").split(",") if f]
SIZES = [int(n) for n in os.environ.get("BVH_SIZES", "1,4,8,16,32").split(",")]
# BVH_LEAVES="4,2,8": triangles per leaf of the host build (RT_BVH_LEAF), first entry = the image the others are compared with
LEAVES = [v for v in os.environ.get("BVH_LEAVES", "").split(",") if v]
dev = G.Device(0)
base = G.init_cornell_box(W, H)
for n in SIZES:
    sc = G.init_cornell_box(W, H)
    if n > 1:
        tv, tm, T = tessellate(base.vertices, base.materials, 36, n)
        sc.vertices, sc.materials, sc.num_triangles = tv, tm, T
    ref = None
    for leaf, form in [(lf, fm) for lf in (LEAVES or [""]) for fm in FORMS]:
        if leaf:
            os.environ["RT_BVH_LEAF"] = leaf
        r = G.Renderer(sc, W, H, device=dev)               # a new queue: a new hierarchy
        os.environ.pop("RT_BVH_FORM", None)
        t0 = time.time(); r.draw(spp=2, variant=L.VARIANT_BVH); build_s = time.time() - t0     # includes the device build
        if form == "-":
            os.environ.pop("RT_BVH_FORM", None)
        else:
            os.environ["RT_BVH_FORM"] = form
        ts = []
        for _ in range(3):
            r.draw_async(spp=SPP, variant=L.VARIANT_BVH).wait_until_completed()
            ts.append(r.last_command.gpu_time_ms())
        ms = min(ts)
        img = r.image()
        if ref is None:
            ref = img
        line = (f"T = {sc.num_triangles:6d} {form:>4s} {('leaf ' + leaf) if leaf else ''}: {ms:8.2f} ms  {W * H * SPP / ms / 1e3:8.1f} Mpaths/s  "
                f"(first draw incl. device build and Python {build_s * 1e3:.0f} ms)  same image as the first form: {np.array_equal(img, ref)}")
        print(line, flush=True)
    os.environ.pop("RT_BVH_FORM", None)
    if n == 1:
        r.draw_async(spp=SPP).wait_until_completed()
        print(f"T = {sc.num_triangles:6d} brute-force default kernel {r.last_command.gpu_time_ms():.2f} ms", flush=True)
