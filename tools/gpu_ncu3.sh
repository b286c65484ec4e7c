#!/bin/bash
# Three `ncu --set full` captures (each after its command ran clean without ncu): bash tools/gpu_ncu3.sh TAG
#   a  the headline kernel, 1080p x 64 spp (profiles/ncu_constants.json comes from this one)
#   b  the ticketed kernel on ONE RANK'S SHARE of the 8-GPU frame (135 interleaved rows of 1080), 1024 spp
#   c  the config-4 kernel (256 spheres + 4 lights), 1080p x 32 spp
TAG=${1:-ncu3}; O=gpurun_out; mkdir -p $O
run() {  # name, then the command
  local name=$1; shift
  timeout 300 "$@" > $O/${TAG}_${name}_plain.log 2>&1 || { echo "$name plain run failed"; tail -5 $O/${TAG}_${name}_plain.log; return; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:pathTrace -s 3 -c 1 -f -o $O/${TAG}_${name} "$@" > $O/${TAG}_${name}_ncu.log 2>&1
  echo "$name ncu exit $?"
}
run a python bench.py --steps 1 --warmup 3 --spp 64 --no-cpu-baseline
run b python tools/rank_share_once.py 8 1024
run c python bench.py --config 4 --steps 1 --warmup 3 --spp 32 --no-cpu-baseline
# launch list of the default bench command (shares of a step, not absolutes)
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > $O/${TAG}_plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file $O/${TAG}_launches.csv $CMD > $O/${TAG}_ncu_launches.log 2>&1
echo "launch list exit $?"; ls -la $O | grep $TAG
