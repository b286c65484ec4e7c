"""Stands in for a racecheck run (compute-sanitizer is closed on this pool): every sorted / ticketed / wide form against the
per-pixel kernel over whole frames, many offset seeds, sample counts, bounce counts and frame shapes, each form TWICE (a race
shows as a frame that differs from its own repeat or from the per-pixel kernel's).  python tools/stress_forms.py [rounds]
-> gpurun_out/stress_forms.json"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import gpuraytracer_b200 as G
from gpuraytracer_b200 import _lib as L

rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 12
rng = np.random.default_rng(2026)
dev = G.Device(0)
forms = [("default", L.VARIANT_DEFAULT), ("sorted", L.VARIANT_SORTED), ("sorted_tiles", L.VARIANT_SORTED_TILES),
         ("sortedblk", L.VARIANT_SORTED_BLOCK), ("sortedblk_tiles", L.VARIANT_SORTED_BLOCK_TILES), ("wave", L.VARIANT_WAVE),
         ("tiles", L.VARIANT_TILES), ("wide", L.VARIANT_WIDE)]
out = dict(frames=0, compared=0, mismatches=[], cases=[])
t0 = time.time()
for k in range(rounds):
    W = int(rng.choice([1920, 1280, 801, 333])); H = int(rng.choice([1080, 720, 451, 97]))
    spp = int(rng.choice([1, 3, 6, 7, 13, 32, 64])); bounces = int(rng.choice([1, 2, 3, 3, 4, 4])); seed = int(rng.integers(1, 1 << 30))
    scene = G.init_cornell_box(W, H) if k % 3 else G.init_sphere_scene(W, H)
    mis = L.VARIANT_MIS if k % 4 == 3 else 0
    r = G.Renderer(scene, W, H, device=dev, seed=seed)
    base = r.draw(spp=spp, bounces=bounces, variant=L.VARIANT_SMEM | mis)
    out["frames"] += 1
    for name, v in forms:
        if mis and name in ("sorted", "sorted_tiles", "tiles", "wide", "sortedblk_tiles"):
            pass                                       # (they fall to the nearest MIS form: still a valid comparison)
        for rep in range(2):
            img = r.draw(spp=spp, bounces=bounces, variant=v | mis)
            out["frames"] += 1; out["compared"] += 1
            if not np.array_equal(img, base):
                out["mismatches"].append(dict(case=k, form=name, rep=rep, W=W, H=H, spp=spp, bounces=bounces, seed=seed,
                                              pixels=int((img != base).any(axis=2).sum())))
    out["cases"].append(dict(W=W, H=H, spp=spp, bounces=bounces, seed=seed, scene="cornell" if k % 3 else "spheres", mis=bool(mis)))
    print(k, W, H, spp, bounces, "spheres" if k % 3 == 0 else "cornell", "mis" if mis else "", len(out["mismatches"]), flush=True)
out["seconds"] = time.time() - t0
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "stress_forms.json"), "w"), indent=1)
print("frames", out["frames"], "compared", out["compared"], "mismatches", len(out["mismatches"]))
