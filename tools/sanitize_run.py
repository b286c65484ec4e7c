"""Small frames through every kernel form, for compute-sanitizer (tools/gpu_sanitize.sh): python tools/sanitize_run.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import gpuraytracer_b200 as G
from gpuraytracer_b200 import _lib as L
from bvh_util import tessellate

W, H = 72, 44                    # ragged against the 16 x 8 block and the 8 x 4 tile
dev = G.Device(0)
r = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev)
first = None
# the ticketed forms cut the samples into chunks only on large launches: force three (6 + 6 + 1 samples) so that
# the chunk hand-over across blocks (wave.cuh: tiles_wait_for_previous_chunk / tiles_publish_chunk) runs here
dev.set_tuning("tiles_chunks", 3)
for name in ("DEFAULT", "SMEM", "CONST", "WAVE", "TILES", "SORTED", "SORTED_TILES", "SORTED_BLOCK", "SORTED_BLOCK_TILES", "BVH"):
    img = r.draw(spp=13, variant=getattr(L, "VARIANT_" + name))
    first = img if first is None else first
    print(f"{name:20s} same image as DEFAULT: {np.array_equal(img, first)}", flush=True)
for spp in (1, 3):               # the wide forms
    a = r.draw(spp=spp); b = r.draw(spp=spp, variant=L.VARIANT_SMEM)
    print(f"wide, {spp} spp            same image as SMEM: {np.array_equal(a, b)}", flush=True)
rs = G.Renderer(G.init_sphere_scene(W, H), W, H, device=dev)
a = rs.draw(spp=3); b = rs.draw(spp=3, variant=L.VARIANT_SMEM)
print(f"spheres                same image as SMEM: {np.array_equal(a, b)}", flush=True)
sc = G.init_cornell_box(W, H)
tv, tm, T = tessellate(sc.vertices, sc.materials, 36, 6)
sc.vertices, sc.materials, sc.num_triangles = tv, tm, T
print("bvh, 1296 triangles    mean", float(G.Renderer(sc, W, H, device=dev).draw(spp=3)[..., :3].mean()), flush=True)
print("rgba8 / rgba16f resolve", r.resolve_rgba8().shape, r.resolve_rgba16f().shape, flush=True)
px, text = G.draw_triangle(G.init_cornell_box(W, H, light_size=1.5), W, H, device=dev, camera_rays=2, mis_samples=30)
print("drawTriangle_mis       mean", float(text.mean()), px.shape, flush=True)
