#!/usr/bin/env python
"""bench.py -- Mpaths/s of the path-tracing kernel on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config 1|2|3|4|5|modeb]
    N > 1:  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One step = one frame of a BASELINE.json workload (default: --config 2 = configs[1], the one the metric is quoted on):
    1      Cornell box 512 x 512, 16 spp                       (configs[0], the CPU-runnable case, here on the GPU)
    2      Cornell box 1920 x 1080, 1024 spp                   (configs[1], bench shape)
    3      Cornell box 3840 x 2160, 4096 spp                   (configs[2], the multi-GPU config)
    4      256 spheres + 4 area lights, 1920 x 1080, 256 spp   (configs[3])
    5      progressive: 1 spp per frame x 1024 frames at 1080p (configs[4]); a step is all 1024 launches
    modeb  drawTriangle (MIS, 6 camera rays x 300 samples) 512 x 512: the "one-bounce MIS" kernel of configs[0]
With N > 1 the SAME frame is dealt to the ranks in interleaved 8-row tiles and gathered to rank 0 over NCCL inside the
timed region (strong scaling: total work fixed).

value : frame paths / device time, inputs resident in HBM, CUDA events on the launch stream.
e2e   : same frame through the public API with HOST buffers: pinned host -> device copies of the scene and the offset
        texture, render, gather, device -> host read of the image.
The JSON line also carries roofline (FP32: algorithmic FLOPs from exact ray counters over the kernel time, against the
nominal AND the measured FFMA peak), cpu_baseline (the C oracle on a bounded row sample, all host cores, N = 1 only)
and parity: the GPU frame -- at N > 1 the GATHERED frame -- against the oracle on the same sampled rows.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Mpaths/s, Cornell 1080p x 1024spp at 1/2/4/8 B200; % of FP32 peak"
UNIT = "Mpaths/s"
TILE_ROWS = 8

CONFIGS = {
    "1": dict(scene="cornell", width=512, height=512, spp=16, frames=1, label="configs[0]"),
    "2": dict(scene="cornell", width=1920, height=1080, spp=1024, frames=1, label="configs[1]"),
    "3": dict(scene="cornell", width=3840, height=2160, spp=4096, frames=1, label="configs[2]"),
    "4": dict(scene="spheres", width=1920, height=1080, spp=256, frames=1, label="configs[3]"),
    "5": dict(scene="cornell", width=1920, height=1080, spp=1, frames=1024, label="configs[4]"),
    "modeb": dict(scene="cornell_modeb", width=512, height=512, spp=6, frames=1, label="configs[0], drawTriangle"),
}


def kernel_source_hash():
    """What the committed ncu figures are tied to: the kernels' sources.  tools/ncu_constants.py writes the same hash into
    profiles/ncu_constants.json; when it differs from the tree's, the figures are from other code and are not reported."""
    h = hashlib.sha256()
    d = os.path.join(ROOT, "gpuraytracer_b200", "csrc")
    for name in sorted(os.listdir(d)):
        if name.endswith((".cu", ".cuh")) or name in ("rt_scene_layout.h", "rt_kernel_smem.h"):
            with open(os.path.join(d, name), "rb") as f:
                h.update(name.encode() + b"\0" + f.read())
    return h.hexdigest()[:16]


def ncu_constants():
    """Figures of the committed `ncu --set full` capture of the headline kernel (tools/ncu_constants.py): DRAM bytes of
    one launch, FP32 flops executed per path, issue-slot use -- or None when the kernels changed since the capture."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_constants.json")) as f:
            nc = json.load(f)
    except (OSError, ValueError):
        return None
    nc["stale"] = nc.get("kernel_source_hash") != kernel_source_hash()
    return nc


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="2", choices=sorted(CONFIGS))
    ap.add_argument("--spp", type=int, default=None)
    ap.add_argument("--bounces", type=int, default=3)
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--frames", type=int, default=None)       # config 5: launches per step
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--width", type=int, default=None)
    ap.add_argument("--height", type=int, default=None)
    a = ap.parse_args()
    c = dict(CONFIGS[a.config])
    custom = False
    for k in ("width", "height", "spp", "frames"):
        v = getattr(a, k)
        if v is not None and v != c[k]:
            c[k] = v
            custom = True
    a.cfg = c
    a.custom = custom
    a.W, a.H, a.spp, a.frames, a.scene = c["width"], c["height"], c["spp"], c["frames"], c["scene"]
    a.modeb = a.scene == "cornell_modeb"
    return a


def workload(a):
    W, H = a.W, a.H
    which = "custom" if a.custom else a.cfg["label"]
    if a.modeb:
        what = f"Cornell box {W}x{H}, drawTriangle: {a.spp} camera rays x 300 MIS samples per pixel"
        scene = "Sources/gpuRaytracer/main.swift scene: 36 triangles, one 1.5 x 1.5 light"
    elif a.scene == "spheres":
        what = f"256 spheres + 4 area lights {W}x{H}, {a.spp} spp, {a.bounces} bounces"
        scene = "SURVEY 8(d) generator, seed 12345: 18 triangles, 256 spheres, 4 lights"
    elif a.frames > 1:
        what = f"Cornell box {W}x{H}, progressive: {a.spp} spp per frame x {a.frames} frames, {a.bounces} bounces"
        scene = "RTrace/scene.swift initCornellBox: 36 triangles, 1 light"
    else:
        what = f"Cornell box {W}x{H}, {a.spp} spp, {a.bounces} bounces"
        scene = "RTrace/scene.swift initCornellBox: 36 triangles, 1 light"
    return {"workload": f"{what} (BASELINE.json {which})", "scene": scene, "width": W, "height": H,
            "spp": a.spp, "frames_per_step": a.frames, "bounces": a.bounces, "offsets": "hash(p + 0x9E3779B9*1) & 0xFFFFF",
            "l2": "inputs are the offsets (4 B/px), the scene (KBs) and the Halton tables (55 MB, read-only, L2-resident); the kernel is "
                  "compute bound (0.02 B/path), so L2 state between steps does not matter; the output is rewritten every step"}


# ------------------------------------------------------------------ clocks ----
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            pass

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def summary(self, t0, t1):
        if self.proc:
            self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1 and len(r) >= 7] or [r for _, r in self.rows[-3:] if len(r) >= 7]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        def num(x):                                        # "[N/A]" on a field must not take the bench down
            try:
                return float(x)
            except ValueError:
                return None
        sm = sorted(v for v in (num(r[0]) for r in rows) if v is not None)
        power = [v for v in (num(r[2]) for r in rows) if v is not None]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(r[3 + k].lower().startswith("active") for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": num(rows[0][1]),
                "power_w_max": max(power) if power else None, "samples": len(rows), "reasons": reasons}


# --------------------------------------------------------------- cpu oracle ----
def sample_tiles(a):
    """Bounded CPU sample of the frame, the SAME rows for the parity check, cpu_baseline and the reference arm:
    7 tiles spread over the image, full width.  Rows per tile are cut so that the sample is 10-30 s of CPU work."""
    H = a.H
    if H <= 512 and not a.modeb:
        return [0], H                                      # config 1: the whole frame is cheap enough
    # CPU cost of a row in units of Cornell paths (14 M/s on 16 threads): a mode-B camera ray is ~410 ray queries, a
    # path through the 256 spheres tests 7 x the primitives
    weight = 80 if a.modeb else 6 if a.scene == "spheres" else 1
    paths_per_row = a.W * a.spp * a.frames * weight
    rows = TILE_ROWS
    while rows > 1 and 7 * rows * paths_per_row > 130e6:
        rows //= 2
    return [int(f * (H - TILE_ROWS)) // TILE_ROWS * TILE_ROWS for f in (0.0, 0.165, 0.336, 0.5, 0.672, 0.836, 1.0)], rows


def oracle_scene(a):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    if a.scene == "spheres":
        return O.spheres(a.W, a.H)
    return O.cornell(a.W, a.H, 1.5 if a.modeb else 1.0)


def oracle_sample(a, nthreads=0):
    """Renders the sample tiles with the C oracle.  Returns ({row0: image rows}, paths, seconds, threads, rows per tile)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    sc = oracle_scene(a)
    tiles, rows = sample_tiles(a)
    out, paths, t = {}, 0, 0.0
    for r0 in tiles:
        t0 = time.perf_counter()
        if a.modeb:
            _, img, cnt = O.drawtriangle(sc, a.W, a.H, row0=r0, rows=rows, camera_rays=a.spp, nthreads=nthreads)
        else:
            # the progressive config accumulates spp x frames samples: the oracle's mean over all of them is what
            # xyz / w of the accumulation buffer must equal
            img, cnt = O.pathtrace(sc, a.W, a.H, a.spp * a.frames, row0=r0, rows=rows, bounces=a.bounces, nthreads=nthreads)
        t += time.perf_counter() - t0
        out[r0] = img
        paths += int(cnt[0])
    return out, paths, t, O.lib.rt_oracle_max_threads(), rows


def run_reference(a, rank):
    """--impl reference: the reference's own kernel source on the host cores, on the same sample tiles the GPU arm's
    cpu_baseline and parity check use.  oracle/_ref is RTrace/raytrace.metal + sampling.metal (and shaders.metal for mode
    B), UNMODIFIED, compiled as C++ (oracle/ref_driver_*.cpp; the closed-source Metal ray query is the brute-force
    intersector of oracle/rt_detmath.h) -- so pathTrace runs at the shader's literal 400 spp / 3 bounces
    (raytrace.metal:24-25); Mpaths/s does not depend on the sample count.  Where the reference has no code for the
    workload (spheres: config 4) or the prebuilt library is missing, the C restatement (oracle port, bit-identical to it
    where both exist) is timed instead.  All host threads, bounded sample per step."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    import ref_lib as R
    W, H = a.W, a.H
    ref = R.load() if (a.bounces == 3 and a.scene != "spheres" and (not a.modeb or a.spp == 6)) else None
    tiles, rows = sample_tiles(a)
    cores = O.lib.rt_oracle_max_threads()
    if ref is not None:
        sc, kind = oracle_scene(a), "reference"
        if a.modeb:
            spp = 6

            def one():
                t0 = time.perf_counter()
                for r0 in tiles:
                    R.drawtriangle(ref, sc, W, H, row0=r0, rows=rows)
                return len(tiles) * rows * W * spp, time.perf_counter() - t0
        else:
            off, spp = O.offsets(W, 0, H, seed=1), 400

            def one():
                t0 = time.perf_counter()
                for r0 in tiles:
                    R.pathtrace(ref, sc, W, H, off, row0=r0, rows=rows)
                return len(tiles) * rows * W * spp, time.perf_counter() - t0
    else:
        spp, kind = a.spp * a.frames, "port"

        def one():
            _, p, dt, _, _ = oracle_sample(a)
            return p, dt
    for _ in range(min(a.warmup, 1)):
        one()
    t, paths = 0.0, 0
    for _ in range(a.steps):
        p, dt = one()
        t += dt; paths += p
    v = paths / t / 1e6
    what = ("oracle/_ref = the reference's shader source compiled as C++, at the shader's literal sample count"
            if kind == "reference" else "C restatement of the shaders (oracle port)")
    sample = (f"{len(tiles)} tiles of {rows} rows x {W} px x {spp} {'camera rays' if a.modeb else 'spp'} per step "
              f"({paths // a.steps} paths), rows {tiles}, {what}, OpenMP {cores} threads")
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus,
                      "steps": a.steps, "warmup": a.warmup, "ms_per_step": t / a.steps * 1e3,
                      "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
                      "data": "synthetic", "config": workload(a),
                      "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
                      "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                      "gpu_launches": 0}))


# ------------------------------------------------------------------ our arm ----
def run_ours(a, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist
    import gpuraytracer_b200 as G
    from gpuraytracer_b200 import _lib as L
    from gpuraytracer_b200 import distributed as D
    from gpuraytracer_b200.api import Buffer, Command, default_params

    W, H = a.W, a.H
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    stream = torch.cuda.Stream(device=dev)
    row0, rows, tile_rows, tile_stride = D.tile_params(H, rank, world, TILE_ROWS)
    pad = D.max_local_rows(H, world, TILE_ROWS)
    my_rows = D.local_rows(H, rank, world, TILE_ROWS)
    progressive = a.frames > 1

    # output rows live in a torch tensor so that NCCL can gather them; everything else is rt* buffers
    if a.modeb:
        scene = G.init_cornell_box(W, H, light_size=1.5)
        rtdev = G.Device(local_rank)
        local = torch.zeros((pad, W, 3), dtype=torch.float32, device=dev)          # the packed float3 text buffer (shaders.metal:627-633)
        pixels = torch.zeros((pad, W, 4), dtype=torch.uint8, device=dev)
        pipeline, queue = rtdev.make_pipeline("drawTriangle"), rtdev.make_queue(stream.cuda_stream)
        bufs = [rtdev.make_buffer(x) for x in (scene.camera, scene.materials, scene.lights, scene.vertices)]
        text_b = Buffer(rtdev, length=rows * W * 12, device_pointer=local.data_ptr())
        pix_b = Buffer(rtdev, length=rows * W * 4, device_pointer=pixels.data_ptr())
        upload = bufs

        def launch():
            cmd = Command(queue, pipeline)
            for b, idx in zip(bufs, (L.INDEX_CAMERA, L.INDEX_MATERIALS, L.INDEX_LIGHTS, L.INDEX_VERTICES)):
                cmd.set_buffer(b, 0, idx)
            cmd.set_buffer(pix_b, 0, L.INDEX_PIXELS)
            cmd.set_buffer(text_b, 0, L.INDEX_DEBUG)
            cmd.set_params(default_params(W, H, row0=row0, rows=rows, numTriangles=scene.num_triangles, numLights=scene.num_lights,
                                          cameraRays=a.spp, misSamples=300, tileRows=tile_rows, tileStride=tile_stride))
            cmd.dispatch_threads((W, rows, 1), (8, 8, 1))
            cmd.commit()
            return [cmd]
        info = rtdev.info()
        fp32_peak = rtdev.measure_fp32_peak()
    else:
        scene = G.init_sphere_scene(W, H) if a.scene == "spheres" else G.init_cornell_box(W, H)
        local = torch.zeros((pad, W, 4), dtype=torch.float32, device=dev)
        offsets_full = G.fill_offsets(W, 0, H, seed=1).reshape(H, W)
        offsets = np.ascontiguousarray(offsets_full[my_rows]).reshape(-1)
        r = G.Renderer(scene, W, H, ordinal=local_rank, row0=row0, rows=rows, offsets=offsets,
                       stream=stream.cuda_stream, output_device_pointer=local.data_ptr())
        over = dict(spp=a.spp, bounces=a.bounces, tileRows=tile_rows, tileStride=tile_stride, variant=a.variant)
        upload = [r.camera_buffer, r.materials_buffer, r.lights_buffer, r.vertex_buffer, r.random_buffer] + \
                 ([r.spheres_buffer] if r.spheres_buffer is not None else [])

        def launch():
            if not progressive:
                return [r.draw_async(**over)]
            with torch.cuda.stream(stream):
                local.zero_()                              # the accumulation target starts from 0 every step
            return [r.draw_async(**{**over, "firstSample": f * a.spp, "accumulate": 1}) for f in range(a.frames)]
        info = r.device.info()
        fp32_peak = r.device.measure_fp32_peak()
    host_img = torch.empty((H, W, local.shape[2]), dtype=torch.float32).pin_memory() if rank == 0 else None

    gather_ev = []                                       # (start, end) CUDA events around this rank's part of the gather

    def step(e2e):
        if e2e:   # host -> device: scene + this rank's offsets, from the buffers' pinned mirrors
            for b in upload:
                b.did_modify()
        cmds = launch()
        with torch.cuda.stream(stream):
            g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            g0.record(stream)
            full = D.gather_rows(local, H, world, rank, TILE_ROWS)
            g1.record(stream)
            gather_ev.append((g0, g1))
            if e2e and rank == 0:
                host_img.copy_(full, non_blocking=True)
        return cmds, full

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for b in upload:
        b.contents()                                     # create the pinned host mirrors once
    for _ in range(a.warmup):
        step(False)
    barrier()

    clocks = ClockSampler(local_rank) if rank == 0 else None
    # ---- device-resident timing: CUDA events on the launching stream ----
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_wall0 = time.time()
    ev0.record(stream)
    cmds = [c for _ in range(a.steps) for c in step(False)[0]]
    ev1.record(stream)
    barrier()
    t_wall1 = time.time()
    ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=dev)
    kernel_ms = sum(c.gpu_time_ms() for c in cmds) / a.steps      # this rank's prep + trace per step
    # where a multi-GPU step goes: every rank's kernel time, and the time between a rank's kernel end and the end of its
    # part of the gather (on rank 0 that includes waiting for the slowest rank)
    gather_ms = sum(g0.elapsed_time(g1) for g0, g1 in gather_ev[-a.steps:]) / a.steps
    per_rank = torch.zeros(2 * world, dtype=torch.float64, device=dev)
    per_rank[2 * rank], per_rank[2 * rank + 1] = kernel_ms, gather_ms
    if world > 1:
        dist.all_reduce(per_rank)
    per_rank = per_rank.tolist()
    launches = sum(c.launch_count() for c in cmds)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    clk = clocks.summary(t_wall0, t_wall1) if clocks else None

    # ---- end to end: host buffers in, host image out ----
    step(True); barrier()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        step(True)
        stream.synchronize()
    barrier()
    e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_s = float(e2e_s.item())
    h2d_t = torch.tensor([sum(b.length for b in upload)], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(h2d_t)

    frame_paths = W * H * a.spp * a.frames
    T, S = scene.num_triangles, scene.num_spheres
    # ---- exact work counters (one extra, untimed launch of the counting variant) ----
    if a.modeb:
        paths = closest = shadow = shaded = None
        flops = None
    else:
        cnt = r.count_rays(**{**over, "spp": a.spp * a.frames})        # (progressive: the same samples in one launch)
        c_t = torch.tensor([cnt["paths"], cnt["closest"], cnt["shadow"], cnt["shaded"]], dtype=torch.int64, device=dev)
        if world > 1:
            dist.all_reduce(c_t)
        paths, closest, shadow, shaded = (int(v) for v in c_t.tolist())
        flops = 45 * T * (closest + shadow) + 18 * S * (closest + shadow) + 200 * shaded + 60 * paths
    _, full = step(False); barrier()                      # restore the frame after the counting launch
    if rank != 0:
        return
    if paths is not None:
        assert paths == frame_paths, (paths, frame_paths)

    nominal = info.smCount * 128 * 2 * (info.clockKHz * 1e3) / 1e12
    my_share = rows / H
    out = {"metric": METRIC, "value": frame_paths * a.steps / (ms * 1e-3) / 1e6, "unit": UNIT, "n_gpus": world,
           "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms / a.steps, "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload(a),
           "clocks": clk,
           "e2e": {"value": frame_paths * a.steps / e2e_s / 1e6, "unit": UNIT,
                   "h2d_bytes_per_step": int(h2d_t.item()), "d2h_bytes_per_step": W * H * 4 * local.shape[2]},
           "gpu_launches": launches * world,
           "ranks": {"kernel_ms": [round(v, 3) for v in per_rank[0::2]], "gather_ms": [round(v, 3) for v in per_rank[1::2]],
                     "note": "per rank and step: prep + trace (CUDA events of the commands); kernel end -> end of the rank's part of the "
                             "gather (rank 0: receive from every rank + index_select into place; includes waiting for the slowest rank)"}}
    nc = ncu_constants()
    headline = not a.modeb and a.scene == "cornell" and (W, H, a.bounces, a.frames) == (1920, 1080, 3, 1)
    if flops is not None:
        achieved = flops * my_share / (kernel_ms * 1e-3) / 1e12
        use_nc = nc if (nc and headline and not nc["stale"]) else None
        exec_flops = use_nc["executed_fp32_flop_per_path"] * paths if use_nc else None
        roof = {"bound": "fp32", "achieved": achieved, "peak": nominal, "unit": "TFLOP/s", "frac": achieved / nominal,
                "peak_kind": f"nominal {info.smCount} SMs x 128 lanes x 2 x {info.clockKHz / 1e3:.0f} MHz (MEASURED_PEAKS.json has no FP32 figure)",
                "peak_measured": fp32_peak[0], "peak_measured_ffma2": fp32_peak[1],
                "peak_measured_kind": "FFMA-only / packed FFMA2-only kernel timed in this run (csrc/pathtrace.cu: rt_fp32_peak)",
                "frac_of_measured": achieved / fp32_peak[0],
                "achieved_note": "ALGORITHMIC flops (SURVEY 8d: every ray charged all T triangles at 45 flop) / kernel time: a culling gain "
                                 "raises it without more arithmetic; executed_* below is what the FP32 pipe does",
                "traffic": use_nc["dram_bytes_per_launch"] if use_nc else None,
                "executed_flops_frac": exec_flops * my_share / (kernel_ms * 1e-3) / 1e12 / nominal if exec_flops else None,
                "executed_flops_per_path": use_nc["executed_fp32_flop_per_path"] if use_nc else None,
                "issue_active_pct": use_nc["issue_active_pct"] if use_nc else None,
                "warp_instructions_per_path": use_nc["warp_instructions_per_path"] if use_nc else None,
                "ncu_capture": (None if not nc else {"source": nc["source"], "kernel": nc["kernel"], "capture": nc["capture"],
                                                     "kernel_source_hash": nc.get("kernel_source_hash"),
                                                     "current_kernel_source_hash": kernel_source_hash(),
                                                     "used": bool(use_nc),
                                                     "why_not": None if use_nc else ("the kernel sources changed since the capture" if nc["stale"] else "captured on the headline workload only")}),
                "kernel": "pathTrace (prep + trace), rank 0", "kernel_ms": kernel_ms,
                "flops_per_frame": flops, "flops_per_path": flops / paths,
                "counters": {"paths": paths, "closest_rays": closest, "shadow_rays": shadow, "shaded_bounces": shaded},
                "hbm_algorithmic_bytes": W * H * (36 if progressive else 20) * a.frames,
                "hbm_gbs": W * H * (36 if progressive else 20) * a.frames * my_share / (kernel_ms * 1e-3) / 1e9}
        if progressive:
            roof["accumulate_pass"] = {"bytes_per_frame": W * H * 36, "what": "4 B offset + 16 B read + 16 B written per pixel and launch",
                                       "ms_per_frame": kernel_ms / a.frames, "hbm_gbs": roof["hbm_gbs"]}
        out["roofline"] = roof
    else:
        out["roofline"] = {"bound": "fp32", "achieved": None, "peak": nominal, "unit": "TFLOP/s", "frac": None,
                           "peak_measured": fp32_peak[0], "kernel": "drawTriangle_mis", "kernel_ms": kernel_ms,
                           "note": "SURVEY 8(d) defines no flop formula for mode B; camera rays / s is the figure"}

    # ---- parity of the (gathered) GPU frame on the sampled rows; CPU baseline at N = 1 ----
    if not a.no_cpu_baseline:
        img = full[:H].cpu().numpy()
        if progressive:
            assert np.all(img[..., 3] == a.spp * a.frames)
            img = img.copy()
            img[..., :3] = img[..., :3] / img[..., 3:4]    # resolve: xyz / w, the oracle's lum / samples
            img[..., 3] = 1.0
        ref, cpaths, ct, cores, tile_rows_cpu = oracle_sample(a)
        bad, equal = 0, True
        for r0, tile in ref.items():
            g = img[r0:r0 + tile_rows_cpu]
            bad += int((np.abs(g - tile) > 1e-4 * np.maximum(np.abs(tile), 1e-6)).sum())
            equal &= bool(np.array_equal(g, tile))
        out["parity"] = {"rows_checked": len(ref) * tile_rows_cpu, "values_out_of_1e-4_rel": bad, "bit_equal": equal,
                         "what": ("gathered frame of %d ranks" % world) if world > 1 else "the frame"}
        if world == 1:
            out["cpu_baseline"] = {"value": cpaths / ct / 1e6, "unit": UNIT, "cores": cores, "kind": "port",
                                   "sample": f"{len(ref)} tiles of {tile_rows_cpu} rows x {W} px x {a.spp * a.frames} "
                                             f"{'camera rays' if a.modeb else 'spp'} ({cpaths} paths, {ct:.1f} s), rows {list(ref)}, OpenMP {cores} threads"}
    print(json.dumps(out))


def main():
    a = parse()
    # stdout carries exactly ONE line (the JSON): libraries that chat on fd 1 (NCCL prints its version there)
    # are sent to stderr, and print() keeps the real stdout
    real = os.fdopen(os.dup(1), "w")
    sys.stdout.flush()
    os.dup2(2, 1)
    sys.stdout = real
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if a.impl == "reference":
        # torchrun exports OMP_NUM_THREADS=1 to its workers when the caller set nothing; the CPU arm is to use all the
        # host's threads (rank 0 alone works), so undo that before libgomp reads it
        if "TORCHELASTIC_RUN_ID" in os.environ and os.environ.get("OMP_NUM_THREADS") == "1":
            os.environ["OMP_NUM_THREADS"] = str(len(os.sched_getaffinity(0)))
        run_reference(a, rank)
        return
    if "TORCHELASTIC_RUN_ID" in os.environ and os.environ.get("OMP_NUM_THREADS") == "1" and rank == 0:
        os.environ["OMP_NUM_THREADS"] = str(len(os.sched_getaffinity(0)))     # rank 0's parity leg runs the oracle
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl")
    try:
        run_ours(a, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
