#!/bin/bash
# kernel iteration: parity tests, variant timings (with and without the Halton tables), bench
TAG=${1:-r04b}
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -15 $O/${TAG}_pytest_gpu.log
timeout 300 python tools/time_variants.py 1920 1080 64 1,3,18,20,21,22 > $O/${TAG}_variants.log 2>&1; cat $O/${TAG}_variants.log
RT_NO_HALTON_TABLE=1 timeout 300 python tools/time_variants.py 1920 1080 64 1,3,20 > $O/${TAG}_variants_notable.log 2>&1; cat $O/${TAG}_variants_notable.log
timeout 600 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench exit $?"; cut -c1-400 $O/${TAG}_bench.json; tail -3 $O/${TAG}_bench.err
