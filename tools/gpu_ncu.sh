#!/bin/bash
# ncu --set full of one pathTrace launch: bash tools/gpu_ncu.sh <variant> <tag> [spp]
V=${1:-0}; TAG=${2:-prof}; SPP=${3:-64}
O=gpurun_out; mkdir -p $O
CMD="python bench.py --steps 1 --warmup 3 --spp $SPP --variant $V --no-cpu-baseline"
timeout 300 $CMD > $O/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pathTrace -s 3 -c 1 \
    -f -o $O/${TAG}_prof $CMD > $O/${TAG}_ncu_full.log 2>&1
echo "ncu exit $?"; tail -2 $O/${TAG}_plain.log | cut -c1-300
