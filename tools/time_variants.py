"""Times kernel variants on one GPU: python tools/time_variants.py [W H SPP] -> table + JSON in gpurun_out/."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import gpuraytracer_b200 as G

W, H, SPP = (int(a) for a in sys.argv[1:4]) if len(sys.argv) >= 4 else (1920, 1080, 64)
variants = [int(v) for v in sys.argv[4].split(",")] if len(sys.argv) > 4 else [1, 2, 3]
dev = G.Device(0)
info = dev.info()
print(info.name.decode(), info.smCount, "SMs", info.clockKHz / 1e3, "MHz")
scene = G.init_sphere_scene(W, H) if os.environ.get('RT_SCENE') == 'spheres' else G.init_cornell_box(W, H)
r = G.Renderer(scene, W, H, device=dev)
res = {}
ref = None
for v in variants:
    r.draw(spp=4, variant=v)                      # warm-up
    ts = []
    for _ in range(3):
        r.draw_async(spp=SPP, variant=v).wait_until_completed()
        ts.append(r.last_command.gpu_time_ms())
    img = r.image()
    if ref is None:
        ref = img
    same = bool(np.array_equal(img, ref))
    ms = min(ts)
    res[v] = dict(ms=ms, all=ts, mpaths=W * H * SPP / ms / 1e3, same_as_first=same)
    print(f"variant {v}: {ms:9.3f} ms  {res[v]['mpaths']:9.1f} Mpaths/s  same={same}  runs={['%.2f' % t for t in ts]}")
os.makedirs("gpurun_out", exist_ok=True)
json.dump(dict(W=W, H=H, spp=SPP, res=res), open("gpurun_out/variants.json", "w"), indent=1)
