#!/bin/bash
# Multi-GPU legs: bash tools/gpu_scale.sh "<N list>" tag   (inside gpurun --gpus max(N))
NS=${1:-2}; TAG=${2:-r01}
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv > $O/${TAG}_smi.csv
for N in $NS; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N \
      bench.py --gpus $N --steps 3 --warmup 3 > $O/${TAG}_bench_n${N}.json 2> $O/${TAG}_bench_n${N}.err
  echo "N=$N exit $?"; cut -c1-260 $O/${TAG}_bench_n${N}.json
done
# BASELINE.json configs[2]: 4K, 4096 spp, row tiles + NCCL gather, on the largest N
N=$(echo $NS | awk '{print $NF}')
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29520 \
    bench.py --gpus $N --steps 2 --warmup 3 --width 3840 --height 2160 --spp 4096 > $O/${TAG}_bench4k_n${N}.json 2> $O/${TAG}_bench4k_n${N}.err
echo "4K N=$N exit $?"; cut -c1-260 $O/${TAG}_bench4k_n${N}.json
