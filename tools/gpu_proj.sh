#!/bin/bash
# parity tests, then stage times of the configs where the projection dominates
set -u
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
for cfg in "c2 256" "c4 512" "c4 1024" "c3 512"; do
  timeout 300 python tools/one_step.py $cfg --ob02 --warm 3 --reps 1 2>&1 | tail -1 | tee -a $O/proj_times.log
done
