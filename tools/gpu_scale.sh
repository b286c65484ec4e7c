#!/bin/bash
# Multi-GPU legs: bash tools/gpu_scale.sh "<N list>" tag ["<configs>"]   (inside gpurun --gpus max(N))
# Every line carries parity of the GATHERED frame against the oracle (bench.py).
NS=${1:-2}; TAG=${2:-r01}; CFGS=${3:-"2 3"}
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv > $O/${TAG}_smi.csv
for c in $CFGS; do
  for N in $NS; do
    steps=3; [ "$c" = "3" ] && steps=2
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N \
        bench.py --gpus $N --steps $steps --warmup 3 --config $c > $O/${TAG}_config_${c}_n${N}.json 2> $O/${TAG}_config_${c}_n${N}.err
    echo "config $c N=$N exit $?"
    python - <<PY
import json
try:
    d = json.load(open("$O/${TAG}_config_${c}_n${N}.json"))
    print("  value %.1f e2e %.1f ms/step %.3f kernel_ms %.3f parity %s" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["roofline"]["kernel_ms"], d.get("parity")))
except Exception as e:
    print("  (no line)", e); print(open("$O/${TAG}_config_${c}_n${N}.err").read()[-1200:])
PY
  done
done
# the same frames through rtGroupRender (one process, peer-to-peer gather behind the C ABI): bash tools/gpu_scale.sh ... with GROUP=1
if [ -n "$GROUP" ]; then
  for c in $CFGS; do for N in $NS; do
    timeout 900 python tools/group_bench.py $N $c 3 > $O/${TAG}_group_${c}_n${N}.json 2> $O/${TAG}_group_${c}_n${N}.err
    echo "group config $c N=$N exit $?"; cut -c1-400 $O/${TAG}_group_${c}_n${N}.json
  done; done
fi
