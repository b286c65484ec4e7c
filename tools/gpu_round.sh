#!/bin/bash
# One gpurun call: GPU parity tests, the bench (both arms), variant timings, then the two ncu passes of
# /opt/skills/guides/B200_PROFILING.md.  Everything lands in gpurun_out/.  Usage (from the repo root):
#   gpurun --timeout 1500 -- 'bash tools/gpu_round.sh [tag]'
TAG=${1:-r01}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $O/${TAG}_smi.csv 2>&1
nproc > $O/${TAG}_nproc.txt
timeout 600 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -3 $O/${TAG}_pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > $O/${TAG}_smoke.log 2>&1; tail -2 $O/${TAG}_smoke.log
timeout 600 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench exit $?"; cat $O/${TAG}_bench.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_reference.json 2>> $O/${TAG}_bench.err; cat $O/${TAG}_bench_reference.json
timeout 300 python tools/time_variants.py 1920 1080 64 1,2 > $O/${TAG}_variants.log 2>&1; cat $O/${TAG}_variants.log
timeout 120 python tools/time_bvh.py > $O/${TAG}_bvh_timing.txt 2>&1; tail -5 $O/${TAG}_bvh_timing.txt
# ncu pass 1: launch list of the bench command (shares, not absolutes)
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > $O/${TAG}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file $O/${TAG}_launches.csv $CMD > $O/${TAG}_ncu_launches.log 2>&1
echo "ncu launches exit $?"
# ncu pass 2: full capture of the top kernel on a short (64 spp) frame
CMD2="python bench.py --steps 1 --warmup 3 --spp 64 --no-cpu-baseline"
timeout 300 $CMD2 > $O/${TAG}_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pathTrace -s 3 -c 1 \
    -f -o $O/${TAG}_prof $CMD2 > $O/${TAG}_ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $O
