#!/bin/bash
# BVH kernel: parity test, then timing on tessellated Cornell scenes (python tools/time_bvh.py)
O=gpurun_out; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "bvh" 2>&1 | tail -15
timeout 300 python tools/time_bvh.py 2>&1 | tee $O/bvh_timing.txt
