#!/bin/bash
# Quick GPU check of the sorted kernel: parity tests, then variant timings for several queue depths R.
#   gpurun --timeout 900 -- 'bash tools/gpu_sorted.sh [tag] [Rs]'
TAG=${1:-s1}; RS=${2:-"6 4 5 8"}
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -25 $O/${TAG}_pytest_gpu.log
for R in $RS; do
  echo "== RT_SORTED_R=$R" | tee -a $O/${TAG}_variants.log
  RT_SORTED_R=$R timeout 300 python tools/time_variants.py 1920 1080 64 1,3,18 2>&1 | tee -a $O/${TAG}_variants.log
done
