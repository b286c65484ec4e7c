#!/bin/bash
# Every BASELINE config as a bench.py workload on ONE GPU: bash tools/gpu_configs.sh TAG [configs]
TAG=${1:-cfg}; CFGS=${2:-"1 2 3 4 5 modeb"}
O=gpurun_out; mkdir -p $O
for c in $CFGS; do
  steps=3; [ "$c" = "3" ] && steps=1
  timeout 900 python bench.py --config $c --steps $steps --warmup 3 > $O/${TAG}_config_$c.json 2> $O/${TAG}_config_$c.err
  echo "config $c exit $?"; python - <<PY
import json
try:
    d = json.load(open("$O/${TAG}_config_$c.json"))
    r = d.get("roofline", {})
    print("  value %.1f e2e %.1f ms/step %.3f launches %s frac %s peak_measured %s parity %s cpu %s" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["gpu_launches"],
          r.get("frac"), r.get("peak_measured"), d.get("parity"), d.get("cpu_baseline", {}).get("value")))
except Exception as e:
    print("  (no line)", e); print(open("$O/${TAG}_config_$c.err").read()[-1500:])
PY
done
