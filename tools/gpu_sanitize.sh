#!/bin/bash
# compute-sanitizer over small frames of every kernel form (memcheck, and racecheck on the shared-memory
# staging, the sort's counters and the tile tickets).  Usage: gpurun --timeout 900 -- 'bash tools/gpu_sanitize.sh [tag]'
TAG=${1:-san}
O=gpurun_out; mkdir -p $O
timeout 120 python tools/sanitize_run.py > $O/${TAG}_plain.log 2>&1 || { tail -5 $O/${TAG}_plain.log; exit 1; }
for tool in memcheck racecheck synccheck; do
    timeout 600 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_run.py > $O/${TAG}_$tool.log 2>&1
    echo "$tool exit $?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard" $O/${TAG}_$tool.log | tail -3
done
