#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q > $O/gpu_all.log 2>&1; tail -3 $O/gpu_all.log | cut -c1-300; grep -E "^FAILED|CUDA error" $O/gpu_all.log | head
for ps in 0 1; do for cfg in "c4 1024" "c3 512" "c2 256"; do echo "point_spec=$ps $cfg"; python tools/one_step.py $cfg --ob02 --reps 3 --gpu point_spec=$ps 2>&1 | tail -1 | cut -c1-400; done; done
