"""profiles/ncu_constants.json from one `ncu --set full` capture of the bench kernel (tools/gpu_round.sh: 64 spp frame):
    python tools/ncu_constants.py gpurun_out/TAG_prof.ncu-rep 1920 1080 64 profiles/TAG_ncu_full_summary.txt > profiles/ncu_constants.json
bench.py reports these per-launch / per-path figures beside the numbers it measures live: DRAM bytes of one launch
(`roofline.traffic`) and the FP32 flops the kernel EXECUTES per path (`roofline.executed_flops_frac`)."""
import csv, io, json, os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import kernel_source_hash       # the figures are tied to the kernel sources they were captured from
rep, W, H, spp, src = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), sys.argv[5]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
d = dict(zip(rows[0], rows[2])); u = dict(zip(rows[0], rows[1]))
f = lambda k: float(d[k].replace(",", ""))
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
dram = f("dram__bytes_read.sum") * scale[u["dram__bytes_read.sum"]] + f("dram__bytes_write.sum") * scale[u["dram__bytes_write.sum"]]
rate = (f("smsp__sass_thread_inst_executed_op_fadd_pred_on.sum.per_cycle_elapsed") + f("smsp__sass_thread_inst_executed_op_fmul_pred_on.sum.per_cycle_elapsed")
        + 2 * f("smsp__sass_thread_inst_executed_op_ffma_pred_on.sum.per_cycle_elapsed"))
flop = rate * f("sm__cycles_elapsed.avg")
print(json.dumps({"kernel": d["Kernel Name"], "capture": f"{W}x{H} x {spp} spp, one launch", "source": src,
                  "kernel_source_hash": kernel_source_hash(),
                  "dram_bytes_per_launch": int(dram), "executed_fp32_flop_per_path": flop / (W * H * spp),
                  "warp_instructions_per_path": f("smsp__inst_executed.sum") / (W * H * spp),
                  "issue_active_pct": f("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                  "lanes_per_instruction": f("smsp__thread_inst_executed_per_inst_executed.ratio")}, indent=1))
