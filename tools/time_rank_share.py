"""Times ONE rank's share of a WORLD-way row-tile split on one GPU (what each GPU does in bench.py --gpus WORLD):
    python tools/time_rank_share.py W H SPP WORLD v1,v2      e.g. 1920 1080 1024 8 3,16
Prints ms, Mpaths/s of the share and the efficiency against the full-frame rate measured in the same run."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import gpuraytracer_b200 as G
from gpuraytracer_b200 import distributed as D

W, H, SPP, WORLD = (int(a) for a in sys.argv[1:5])
variants = [int(v) for v in sys.argv[5].split(",")]
dev = G.Device(0)
scene = G.init_cornell_box(W, H)
offsets_full = G.fill_offsets(W, 0, H, seed=1).reshape(H, W)


def best(r, **over):
    r.draw(**dict(over, spp=8))
    ts = []
    for _ in range(3):
        r.draw_async(**over).wait_until_completed()
        ts.append(r.last_command.gpu_time_ms())
    return min(ts)


for v in variants:
    full = G.Renderer(scene, W, H, device=dev)
    ms_full = best(full, spp=SPP, variant=v)
    img_full = full.image()
    line = f"variant {v}: full {ms_full:8.2f} ms {W * H * SPP / ms_full / 1e3:7.1f} Mpaths/s |"
    for rank in (0, WORLD - 1):
        row0, rows, tr, tstride = D.tile_params(H, rank, WORLD, 8)
        mine = D.local_rows(H, rank, WORLD, 8)
        off = np.ascontiguousarray(offsets_full[mine]).reshape(-1)
        r = G.Renderer(scene, W, H, device=dev, row0=row0, rows=rows, offsets=off)
        ms = best(r, spp=SPP, variant=v, tileRows=tr, tileStride=tstride)
        same = bool(np.array_equal(r.image(), img_full[mine]))
        line += f" rank {rank}: {ms:7.2f} ms, x{WORLD} = {ms_full / ms:5.2f} ({ms_full / ms / WORLD * 100:5.1f} %) same={same} |"
    print(line, flush=True)
