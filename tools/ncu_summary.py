"""Summarise an .ncu-rep (read here, no GPU): python tools/ncu_summary.py rep.ncu-rep [out.txt]
Prints, per captured launch, the counters DESIGN.md argues from: duration, registers, occupancy,
issue-slot and pipe utilisation, executed FP32 instruction counts, DRAM bytes, top stall reasons."""
import csv, io, subprocess, sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
WANT = [
    "Kernel Name", "Block Size", "Grid Size", "gpu__time_duration.sum", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__waves_per_multiprocessor",
    "smsp__cycles_active.avg", "sm__cycles_elapsed.avg", "sm__cycles_elapsed.avg.per_second",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed.avg.per_cycle_elapsed", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fmalite_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_cbu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_adu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum", "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum",
    "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum", "smsp__sass_thread_inst_executed_op_fp32_pred_on.sum",
    "smsp__sass_thread_inst_executed_op_integer_pred_on.sum", "smsp__sass_thread_inst_executed_op_control_pred_on.sum",
    "smsp__sass_thread_inst_executed_op_conversion_pred_on.sum", "smsp__sass_thread_inst_executed_op_memory_pred_on.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__warps_eligible.avg.per_cycle_active", "smsp__warps_active.avg.per_cycle_active",
]
out = []
for r in rows[2:]:
    d = dict(zip(hdr, r))
    out.append("=" * 100)
    for k in WANT:
        if k in d:
            out.append(f"{k:90s} {units[hdr.index(k)]:14s} {d[k]}")
    # executed FP32 flops: thread-level FADD + FMUL + 2 FFMA, from the per-cycle rates (summed over sub-partitions) x cycles
    try:
        f = lambda k: float(d[k].replace(",", ""))
        rate = (f("smsp__sass_thread_inst_executed_op_fadd_pred_on.sum.per_cycle_elapsed") + f("smsp__sass_thread_inst_executed_op_fmul_pred_on.sum.per_cycle_elapsed")
                + 2 * f("smsp__sass_thread_inst_executed_op_ffma_pred_on.sum.per_cycle_elapsed"))
        cycles = f("sm__cycles_elapsed.avg")
        peak = f("sm__sass_thread_inst_executed_op_ffma_pred_on.sum.peak_sustained") * 2
        out.append(f"{'executed FP32 flop (FADD + FMUL + 2 FFMA, thread level)':90s} {'flop':14s} {rate * cycles:.6e}")
        out.append(f"{'executed FP32 flop per cycle / peak (SMs x 128 x 2)':90s} {'':14s} {rate:.1f} / {peak:.0f} = {rate / peak * 100:.2f} %")
    except (KeyError, ValueError):
        pass
    stalls = [(float(d[h].replace(",", "") or 0), h) for h in hdr
              if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio")]
    out.append("-- warp stall reasons (avg warps stalled per issue-active cycle), top 8")
    for v, h in sorted(stalls, reverse=True)[:8]:
        out.append(f"   {h:88s} {v:.3f}")
text = "\n".join(out)
print(text)
if len(sys.argv) > 2:
    open(sys.argv[2], "w").write(text + "\n")
