"""Render the bench frame with the default (wave) kernel and with the per-pixel kernel (no candidate lists, no
compaction), list the pixels where they differ, and ask the oracle about those rows."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import gpuraytracer_b200 as G
import oracle_lib as O
W, H, SPP = 1920, 1080, int(sys.argv[1]) if len(sys.argv) > 1 else 1024
dev = G.Device(0)
r = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev)
a = r.draw(spp=SPP, variant=3).copy()
b = r.draw(spp=SPP, variant=1).copy()
ys, xs = np.nonzero((a != b).any(-1))
print("wave vs smem differing pixels:", len(ys), list(zip(ys.tolist(), xs.tolist()))[:20])
osc = O.cornell(W, H)
for y in sorted(set(ys.tolist()))[:6]:
    ref, _ = O.pathtrace(osc, W, H, SPP, row0=y, rows=1)
    for x in xs[ys == y].tolist():
        print(f"pixel ({y},{x}): wave {a[y, x]} smem {b[y, x]} oracle {ref[0, x]}  wave==oracle {np.array_equal(a[y,x], ref[0,x])} smem==oracle {np.array_equal(b[y,x], ref[0,x])}")
if len(ys) == 0:
    # both agree: compare a band of rows against the oracle
    for y in (536, 537, 538, 539, 540, 541, 542, 543):
        ref, _ = O.pathtrace(osc, W, H, SPP, row0=y, rows=1)
        bad = np.nonzero((a[y] != ref[0]).any(-1))[0]
        print("row", y, "differs from oracle at x =", bad.tolist()[:10])
        for x in bad.tolist()[:3]:
            print("   gpu", a[y, x], "oracle", ref[0, x])
