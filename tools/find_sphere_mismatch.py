"""Where do two kernel forms differ on the sphere scene?  python tools/find_sphere_mismatch.py [spp]  -> gpurun_out/sphere_mismatch.json"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import gpuraytracer_b200 as G
from gpuraytracer_b200 import _lib as L
import oracle_lib as O
W, H = 1920, 1080
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 4
r = G.Renderer(G.init_sphere_scene(W, H), W, H)
out = {}
for bounces in (1, 2, 3):
    a = r.draw(spp=spp, bounces=bounces, variant=L.VARIANT_SMEM)
    b = r.draw(spp=spp, bounces=bounces, variant=L.VARIANT_WAVE)
    bad = np.argwhere((a != b).any(axis=2))
    rows = sorted(set(int(y) for y, x in bad))[:3]
    rec = dict(count=int(len(bad)), first=[[int(y), int(x), a[y, x].tolist(), b[y, x].tolist()] for y, x in bad[:6]], oracle=[])
    osc = O.spheres(W, H)
    for y in rows:
        ref, _ = O.pathtrace(osc, W, H, spp, row0=y, rows=1, bounces=bounces)
        xs = [int(x) for yy, x in bad if yy == y][:4]
        rec["oracle"].append([y, [[x, ref[0, x].tolist(), bool(np.array_equal(ref[0, x], a[y, x])), bool(np.array_equal(ref[0, x], b[y, x]))] for x in xs]])
    out[bounces] = rec
    print(bounces, rec["count"], flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/sphere_mismatch.json", "w"), indent=1)
