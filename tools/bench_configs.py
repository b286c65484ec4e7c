"""Timings of the BASELINE.json configs that bench.py does not headline, on ONE GPU:
    python tools/bench_configs.py [--quick] > gpurun_out/configs.jsonl
  config 3 shape  Cornell 3840x2160, 4096 spp (single GPU leg; the multi-GPU leg is bench.py --width/--height)
  config 4        256 spheres + 4 area lights, 1920x1080, 256 spp
  config 5        progressive: 1 spp/frame x 1024 frames at 1080p (launch overhead, accumulate GB/s)
  mode B          drawTriangle (MIS) 512x512 and 800x600 at the shader's literals (6 x 300)
Each line: device time from the command's CUDA events, Mpaths/s, algorithmic TFLOP/s with the exact
counters, and a parity check of a few rows against the oracle (the checker, never the thing timed)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import gpuraytracer_b200 as G  # noqa: E402
import oracle_lib as O  # noqa: E402

quick = "--quick" in sys.argv
dev = G.Device(0)
info = dev.info()
PEAK = info.smCount * 128 * 2 * info.clockKHz * 1e3 / 1e12


def emit(**kw):
    print(json.dumps(kw), flush=True)


def flops(c, T, S):
    return 45 * T * (c["closest"] + c["shadow"]) + 18 * S * (c["closest"] + c["shadow"]) + 200 * c["shaded"] + 60 * c["paths"]


def parity_rows(r_scene, o_scene, W, H, spp, rows, img):
    bad = 0
    for r0 in rows:
        ref, _ = O.pathtrace(o_scene, W, H, spp, row0=r0, rows=2)
        bad += int((img[r0:r0 + 2] != ref).sum())
    return bad


def time_frame(r, n=3, **over):
    r.draw(spp=min(over.get("spp", 4), 4), **{k: v for k, v in over.items() if k != "spp"})      # warm-up
    ts = []
    for _ in range(n):
        r.draw_async(**over).wait_until_completed()
        ts.append(r.last_command.gpu_time_ms())
    return min(ts), ts


# ---- config 3 shape on one GPU -----------------------------------------------------------
W, H, SPP = 3840, 2160, (256 if quick else 4096)
r = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev)
ms, ts = time_frame(r, n=1 if not quick else 2, spp=SPP)
img = r.image()
bad = parity_rows(None, O.cornell(W, H), W, H, SPP, [1000], img) if quick else None
emit(config="3: Cornell 3840x2160 (1 GPU leg)", spp=SPP, ms=ms, mpaths_s=W * H * SPP / ms / 1e3, runs=ts,
     values_differing_on_checked_rows=bad)
del r

# ---- config 4: spheres ----------------------------------------------------------------------
W, H, SPP = 1920, 1080, (32 if quick else 256)
sc = G.init_sphere_scene(W, H)
r = G.Renderer(sc, W, H, device=dev)
ms, ts = time_frame(r, spp=SPP)
img = r.image()
c = r.count_rays(spp=SPP)
F = flops(c, sc.num_triangles, sc.num_spheres)
t0 = time.perf_counter()
ref, ocnt = O.pathtrace(O.spheres(W, H), W, H, SPP, row0=540, rows=2)
cpu_s = time.perf_counter() - t0
emit(config="4: 256 spheres + 4 area lights 1920x1080", spp=SPP, ms=ms, mpaths_s=W * H * SPP / ms / 1e3, runs=ts,
     counters=c, flops_per_path=F / c["paths"], tflops=F / ms / 1e9, frac_fp32_peak=F / ms / 1e9 / PEAK,
     values_differing_on_checked_rows=int((img[540:542] != ref).sum()),
     cpu_oracle_mpaths_s=int(ocnt[0]) / cpu_s / 1e6, cpu_threads=O.lib.rt_oracle_max_threads())
del r

# ---- config 5: progressive accumulation -----------------------------------------------------
W, H, FRAMES = 1920, 1080, (64 if quick else 1024)
r = G.Renderer(G.init_cornell_box(W, H), W, H, device=dev)
r.clear()
r.draw_async(spp=1, firstSample=0, accumulate=1).wait_until_completed()
r.clear()
t0 = time.perf_counter()
cmds = [r.draw_async(spp=1, firstSample=f, accumulate=1) for f in range(FRAMES)]
cmds[-1].wait_until_completed()
wall = time.perf_counter() - t0
kernel_ms = sum(cm.gpu_time_ms() for cm in cmds)
acc = r.image()
one, _ = time_frame(r, spp=FRAMES)
single = r.image()
same = bool(np.array_equal(acc[..., :3] / np.float32(FRAMES), single[..., :3]))
bytes_per_frame = W * H * 36                       # 4 B offset + 16 B read + 16 B write per pixel
emit(config="5: progressive 1 spp x frames, 1080p", frames=FRAMES, wall_ms=wall * 1e3, sum_command_ms=kernel_ms,
     ms_per_frame_wall=wall * 1e3 / FRAMES, ms_per_frame_device=kernel_ms / FRAMES, mpaths_s=W * H * FRAMES / wall / 1e6,
     one_launch_same_spp_ms=one, overhead_vs_one_launch=wall * 1e3 / one,
     accumulate_hbm_gbs=bytes_per_frame / (kernel_ms / FRAMES * 1e-3) / 1e9,
     launches=sum(cm.launch_count() for cm in cmds), bit_equal_to_one_launch=same)
del r, cmds

# ---- mode B ----------------------------------------------------------------------------------
for (W, H) in ((512, 512), (800, 600)):
    scb = G.init_cornell_box(W, H, light_size=1.5)
    G.draw_triangle(scb, W, H, device=dev, camera_rays=1, mis_samples=30)
    best = None
    for _ in range(2):
        px, tx, cmd = G.draw_triangle(scb, W, H, device=dev, return_command=True)
        best = cmd.gpu_time_ms() if best is None else min(best, cmd.gpu_time_ms())
    t0 = time.perf_counter()
    wpx, wtx, cnt = O.drawtriangle(O.cornell(W, H, 1.5), W, H, row0=H // 2, rows=4)
    cpu_s = time.perf_counter() - t0
    emit(config=f"mode B drawTriangle (MIS 6 x 300) {W}x{H}", ms=best, camera_rays_per_s=W * H * 6 / best / 1e3,
         values_differing_on_checked_rows=int((tx[H // 2:H // 2 + 4] != wtx).sum()),
         cpu_oracle_camera_rays_per_s=int(cnt[0]) / cpu_s, cpu_threads=O.lib.rt_oracle_max_threads(),
         rays_per_camera_ray_on_checked_rows=int(cnt[1]) / max(int(cnt[0]), 1))
