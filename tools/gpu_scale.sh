#!/bin/bash
# Multi-GPU leg: bash tools/gpu_scale.sh N tag   (inside gpurun --gpus N)
N=${1:-2}; TAG=${2:-r01}
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv > $O/${TAG}_smi_${N}.csv
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 3 --warmup 3 > $O/${TAG}_bench_n${N}.json 2> $O/${TAG}_bench_n${N}.err
echo "exit $?"; cat $O/${TAG}_bench_n${N}.json; tail -5 $O/${TAG}_bench_n${N}.err
