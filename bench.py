#!/usr/bin/env python
"""bench.py -- Mpaths/s of the path-tracing kernel on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    N > 1:  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One step = one frame of the workload BASELINE.json's metric is quoted on:
Cornell box 1920x1080, 1024 spp, 3 bounces (configs[1]).  With N > 1 the SAME frame is
dealt to the ranks in interleaved 8-row tiles and gathered to rank 0 over NCCL inside
the timed region (strong scaling: total work fixed).

value : frame paths / device time, inputs resident in HBM, CUDA events on the launch stream.
e2e   : same frame through the public API with HOST buffers: pinned host -> device copies of
        the scene and the offset texture, render, gather, device -> host read of the image.
The JSON line also carries roofline (FP32, algorithmic FLOPs from exact ray counters),
cpu_baseline (the C oracle on a bounded row sample, all host cores) and the parity of the
GPU frame against that sample.
"""
import argparse
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
W, H = 1920, 1080
TILE_ROWS = 8
def ncu_constants():
    """Figures taken from the committed `ncu --set full` capture of this kernel (tools/ncu_constants.py writes the
    file from the .ncu-rep): DRAM bytes of one launch and the FP32 flops the kernel executes per path."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_constants.json")) as f:
            return json.load(f)
    except (OSError, ValueError):
        return None


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--spp", type=int, default=1024)
    ap.add_argument("--bounces", type=int, default=3)
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--width", type=int, default=W)       # BASELINE.json configs[2]: --width 3840 --height 2160 --spp 4096
    ap.add_argument("--height", type=int, default=H)
    a = ap.parse_args()
    globals()["W"], globals()["H"] = a.width, a.height
    return a


def workload(a):
    which = "configs[1]" if (W, H, a.spp) == (1920, 1080, 1024) else "configs[2]" if (W, H, a.spp) == (3840, 2160, 4096) else "custom"
    return {"workload": f"Cornell box {W}x{H}, {a.spp} spp, {a.bounces} bounces (BASELINE.json {which})",
            "scene": "RTrace/scene.swift initCornellBox: 36 triangles, 1 light", "width": W, "height": H,
            "spp": a.spp, "bounces": a.bounces, "offsets": "hash(p + 0x9E3779B9*1) & 0xFFFFF",
            "l2": "inputs are 8.3 MB offsets + 3.6 KB scene; the kernel is compute bound (0.02 B/path), "
                  "so L2 state between steps does not matter; the 33 MB output is rewritten every step"}


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
def sample_rows():
    """Bounded CPU sample of the frame: 7 tiles of 8 rows spread over the image (56 rows, full width)."""
    return [int(f * (H - TILE_ROWS)) // TILE_ROWS * TILE_ROWS for f in (0.0, 0.165, 0.336, 0.5, 0.672, 0.836, 1.0)]


def oracle_sample(a, rows_per_tile=TILE_ROWS, tiles=None, nthreads=0):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np
    import oracle_lib as O
    sc = O.cornell(W, H)
    tiles = sample_rows() if tiles is None else tiles
    out, paths, t = {}, 0, 0.0
    for r0 in tiles:
        t0 = time.perf_counter()
        img, cnt = O.pathtrace(sc, W, H, a.spp, row0=r0, rows=rows_per_tile, bounces=a.bounces, nthreads=nthreads)
        t += time.perf_counter() - t0
        out[r0] = img
        paths += int(cnt[0])
    return out, paths, t, O.lib.rt_oracle_max_threads()


def run_reference(a, rank):
    """--impl reference: the reference's own kernel source on the host cores.  oracle/_ref is RTrace/raytrace.metal +
    sampling.metal, UNMODIFIED, compiled as C++ (oracle/ref_driver_a.cpp; the closed-source Metal ray query is the
    brute-force intersector of oracle/rt_detmath.h) -- so it runs at the shader's literal 400 spp / 3 bounces
    (raytrace.metal:24-25); Mpaths/s does not depend on the sample count.  Without the prebuilt library the C
    restatement (oracle port, bit-identical to it) is timed instead.  All host threads, bounded sample per step."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    import ref_lib as R
    ref = R.load() if a.bounces == 3 else None
    r0 = H // 2 // TILE_ROWS * TILE_ROWS               # 8 rows through the middle of the frame
    cores = O.lib.rt_oracle_max_threads()
    if ref is not None:
        sc, off, spp, kind = O.cornell(W, H), O.offsets(W, 0, H, seed=1), 400, "reference"

        def one():
            t0 = time.perf_counter()
            R.pathtrace(ref, sc, W, H, off, row0=r0, rows=TILE_ROWS)
            return W * TILE_ROWS * spp, time.perf_counter() - t0
    else:
        spp, kind = a.spp, "port"

        def one():
            _, p, dt, _ = oracle_sample(a, tiles=[r0])
            return p, dt
    for _ in range(a.warmup):
        one()
    t, paths = 0.0, 0
    for _ in range(a.steps):
        p, dt = one()
        t += dt; paths += p
    v = paths / t / 1e6
    what = ("oracle/_ref = the reference's raytrace.metal + sampling.metal compiled as C++, at the shader's literal 400 spp"
            if kind == "reference" else "C restatement of the shaders (oracle port)")
    sample = f"rows {r0}..{r0 + 7} x {W} px x {spp} spp per step ({paths // a.steps} paths), {what}, OpenMP {cores} threads"
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
    from gpuraytracer_b200 import distributed as D

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    stream = torch.cuda.Stream(device=dev)
    row0, rows, tile_rows, tile_stride = D.tile_params(H, rank, world, TILE_ROWS)
    pad = D.max_local_rows(H, world, TILE_ROWS)
    my_rows = D.local_rows(H, rank, world, TILE_ROWS)

    # output rows live in a torch tensor so that NCCL can gather them; everything else is rt* buffers
    local = torch.zeros((pad, W, 4), dtype=torch.float32, device=dev)
    offsets_full = G.fill_offsets(W, 0, H, seed=1).reshape(H, W)
    offsets = np.ascontiguousarray(offsets_full[my_rows]).reshape(-1)
    scene = G.init_cornell_box(W, H)
    r = G.Renderer(scene, W, H, ordinal=local_rank, row0=row0, rows=rows, offsets=offsets,
                   stream=stream.cuda_stream, output_device_pointer=local.data_ptr())
    over = dict(spp=a.spp, bounces=a.bounces, tileRows=tile_rows, tileStride=tile_stride, variant=a.variant)
    host_img = torch.empty((H, W, 4), dtype=torch.float32).pin_memory() if rank == 0 else None
    kernel_ms = []

    def step(e2e):
        if e2e:   # host -> device: scene + this rank's offsets, from the buffers' pinned mirrors
            for b in (r.camera_buffer, r.materials_buffer, r.lights_buffer, r.vertex_buffer, r.random_buffer):
                b.did_modify()
        cmd = r.draw_async(**over)
        with torch.cuda.stream(stream):
            full = D.gather_rows(local, H, world, rank, TILE_ROWS)
            if e2e and rank == 0:
                host_img.copy_(full, non_blocking=True)
        return cmd, full

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for b in (r.camera_buffer, r.materials_buffer, r.lights_buffer, r.vertex_buffer, r.random_buffer):
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
    cmds = [step(False)[0] for _ in range(a.steps)]
    ev1.record(stream)
    barrier()
    t_wall1 = time.time()
    ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=dev)
    kernel_ms = [c.gpu_time_ms() for c in cmds]
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
    h2d = sum(b.length for b in (r.camera_buffer, r.materials_buffer, r.lights_buffer, r.vertex_buffer, r.random_buffer))
    h2d_t = torch.tensor([h2d], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(h2d_t)

    # ---- exact work counters (one extra, untimed launch of the counting variant) ----
    cnt = r.count_rays(**over)
    c_t = torch.tensor([cnt["paths"], cnt["closest"], cnt["shadow"], cnt["shaded"]], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(c_t)
    paths, closest, shadow, shaded = (int(v) for v in c_t.tolist())
    step(False); barrier()                               # restore the frame after the counting launch

    if rank != 0:
        return
    frame_paths = W * H * a.spp
    assert paths == frame_paths, (paths, frame_paths)
    T = scene.num_triangles
    flops = 45 * T * (closest + shadow) + 200 * shaded + 60 * paths          # the contract figure
    # what the kernel EXECUTES (it shares work between the tests of a query and skips camera-ray tests the formula
    # charges): thread-level FADD + FMUL + 2 FFMA per path from the committed ncu capture, or null without one
    nc = ncu_constants()
    flops_exec = nc["executed_fp32_flop_per_path"] * paths if nc and a.width == 1920 and a.height == 1080 and a.bounces == 3 else None
    info = r.device.info()
    k_ms = sum(kernel_ms) / len(kernel_ms)                # this rank's prep+trace per step
    peak = info.smCount * 128 * 2 * (info.clockKHz * 1e3) / 1e12
    my_share = rows / H
    achieved = flops * my_share / (k_ms * 1e-3) / 1e12
    out = {"metric": METRIC, "value": frame_paths * a.steps / (ms * 1e-3) / 1e6, "unit": UNIT, "n_gpus": world,
           "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms / a.steps, "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload(a),
           "clocks": clk,
           "e2e": {"value": frame_paths * a.steps / e2e_s / 1e6, "unit": UNIT,
                   "h2d_bytes_per_step": int(h2d_t.item()), "d2h_bytes_per_step": W * H * 16},
           "gpu_launches": launches * world,
           "roofline": {"bound": "fp32", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                        "traffic": nc["dram_bytes_per_launch"] if nc else None,
                        "traffic_note": "dram__bytes_read+write of one 1080p pathTrace launch from the ncu --set full capture "
                                        + (nc["source"] if nc else "(none)") +
                                        " (offsets read once; the 33 MB of output is still in L2 when the launch ends)",
                        "executed_flops_frac": flops_exec * my_share / (k_ms * 1e-3) / 1e12 / peak if flops_exec else None,
                        "executed_flops_per_path": flops_exec / paths if flops_exec else None,
                        "executed_flops_note": "thread-level FADD + FMUL + 2 FFMA of the same ncu capture, per path",
                        "peak_kind": f"nominal {info.smCount} SMs x 128 lanes x 2 x {info.clockKHz / 1e3:.0f} MHz "
                                     "(MEASURED_PEAKS.json has no FP32 figure)",
                        "kernel": "pathTrace (prep + trace), rank 0", "kernel_ms": k_ms,
                        "flops_per_frame": flops, "flops_per_path": flops / paths,
                        "counters": {"paths": paths, "closest_rays": closest, "shadow_rays": shadow, "shaded_bounces": shaded},
                        "hbm_algorithmic_bytes": W * H * 20, "hbm_gbs": W * H * 20 * my_share / (k_ms * 1e-3) / 1e9}}

    # ---- CPU baseline + parity of the GPU frame on the sampled rows ----
    if world == 1 and not a.no_cpu_baseline:
        img = local[:H].cpu().numpy()
        ref, cpaths, ct, cores = oracle_sample(a)
        bad, equal = 0, True
        for r0, tile in ref.items():
            g = img[r0:r0 + TILE_ROWS]
            bad += int((np.abs(g - tile) > 1e-4 * np.maximum(np.abs(tile), 1e-6)).sum())
            equal &= bool(np.array_equal(g, tile))
        out["cpu_baseline"] = {"value": cpaths / ct / 1e6, "unit": UNIT, "cores": cores, "kind": "port",
                               "sample": f"{len(ref)} tiles of {TILE_ROWS} rows x {W} px x {a.spp} spp "
                                         f"({cpaths} paths, {ct:.1f} s), OpenMP {cores} threads"}
        out["parity"] = {"rows_checked": len(ref) * TILE_ROWS, "values_out_of_1e-4_rel": bad, "bit_equal": equal}
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
