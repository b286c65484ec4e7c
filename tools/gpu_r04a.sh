#!/bin/bash
# round-2 first GPU call: compute-sanitizer over every kernel form, then the parity tests and one bench of the unchanged kernels
TAG=${1:-r04a}
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $O/${TAG}_smi.csv 2>&1
nproc > $O/${TAG}_nproc.txt
bash tools/gpu_sanitize.sh ${TAG}_san
timeout 600 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -3 $O/${TAG}_pytest_gpu.log
timeout 600 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench exit $?"; cut -c1-600 $O/${TAG}_bench.json
