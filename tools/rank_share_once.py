"""One rank's share of an N-GPU frame rendered alone on one GPU, four launches (for ncu: -s 3 -c 1 takes the fourth):
    python tools/rank_share_once.py N [spp]      (rank 0 of N, 1920 x 1080, interleaved 8-row tiles)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import gpuraytracer_b200 as G
from gpuraytracer_b200 import distributed as D
N = int(sys.argv[1]) if len(sys.argv) > 1 else 8
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
W, H = 1920, 1080
row0, rows, tr, ts = D.tile_params(H, 0, N, 8)
mine = D.local_rows(H, 0, N, 8)
off = np.ascontiguousarray(G.fill_offsets(W, 0, H, seed=1).reshape(H, W)[mine]).reshape(-1)
r = G.Renderer(G.init_cornell_box(W, H), W, H, row0=row0, rows=rows, offsets=off)
for k in range(4):
    r.draw_async(spp=spp, tileRows=tr, tileStride=ts).wait_until_completed()
    print(f"launch {k}: {r.last_command.gpu_time_ms():.3f} ms, {rows} rows, ideal share of a {W}x{H} frame: 1/{N}", flush=True)
