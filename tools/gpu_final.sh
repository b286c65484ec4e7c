#!/bin/bash
# The round's last measurement pass on ONE GPU: bash tools/gpu_final.sh TAG
#   GPU tests, smoke, every config, `ncu --set full` of the headline / config-4 / mode-B kernels (summarised ON the box:
#   the .ncu-rep files stay in /tmp there, gpurun_out carries 64 MiB at most), ncu constants, launch list.
TAG=${1:-final}; O=gpurun_out; mkdir -p $O /tmp/rep
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; tail -1 $O/${TAG}_pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1 | tee $O/${TAG}_smoke.log
bash tools/gpu_configs.sh $TAG "1 2 3 4 5 modeb" > $O/${TAG}_configs.log 2>&1; grep -c "exit 0" $O/${TAG}_configs.log
cap() {  # name, kernel regex, then the command
  local name=$1 rx=$2; shift 2
  timeout 300 "$@" > $O/${TAG}_${name}_plain.log 2>&1 || { echo "$name plain run failed"; return; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$rx -s 3 -c 1 -f -o /tmp/rep/$name "$@" > $O/${TAG}_${name}_ncu.log 2>&1
  python tools/ncu_summary.py /tmp/rep/$name.ncu-rep $O/${TAG}_${name}_ncu_full_summary.txt > /dev/null
  echo "$name ncu exit $?"
}
cap a pathTrace python bench.py --steps 1 --warmup 3 --spp 64 --no-cpu-baseline
python tools/ncu_constants.py /tmp/rep/a.ncu-rep 1920 1080 64 profiles/${TAG}_a_ncu_full_summary.txt > $O/${TAG}_ncu_constants.json
B=$(cuobjdump -sass -fun pathTrace_sortedblk6 gpuraytracer_b200/csrc/build/pathtrace.cubin | grep BAR.SYNC | sed 's/.*\/\*\([0-9a-f]*\)\*\/.*/0x\1/' | tr '\n' ' ')
set -- $B; echo "barriers: $B" > $O/${TAG}_regions.txt
# barriers of the entry: staging | tables | camera | pass A | sort (read ranks | write places) | pass B | (fallback body follows)
python tools/ncu_regions.py /tmp/rep/a.ncu-rep 0x0:$2:stage $2:$3:camera $3:$4:passA $4:$6:sort $6:$7:passB $7:$8:accum >> $O/${TAG}_regions.txt
cap c pathTrace python bench.py --config 4 --steps 1 --warmup 3 --spp 32 --no-cpu-baseline
cap modeb drawTriangle python bench.py --config modeb --steps 1 --warmup 3 --no-cpu-baseline
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > $O/${TAG}_plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file $O/${TAG}_launches.csv $CMD > $O/${TAG}_ncu_launches.log 2>&1
echo "launch list exit $?"; du -sh $O
