#!/bin/bash
# A/B timing of experimental library builds (tools/build_exp.sh): bash tools/gpu_ab.sh TAG "name1 name2 ..." [variants] [spp]
TAG=${1:-ab}; NAMES=${2:-""}; VARS=${3:-20}; SPP=${4:-64}
O=gpurun_out; mkdir -p $O
echo "== default" | tee $O/${TAG}_ab.log
timeout 300 python tools/time_variants.py 1920 1080 $SPP 1,$VARS 2>&1 | tee -a $O/${TAG}_ab.log
for n in $NAMES; do
  echo "== $n" | tee -a $O/${TAG}_ab.log
  RTCUDA_LIB=tools/_build/librtcuda_$n.so timeout 300 python tools/time_variants.py 1920 1080 $SPP 1,$VARS 2>&1 | tee -a $O/${TAG}_ab.log
done
