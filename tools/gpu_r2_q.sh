#!/bin/bash
# round 2, call q: opcode carried by the list entries, unit slot hoisted out of the bisection loop
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
for cfg in "c4 1024" "c3 512" "c2 256" "c4 2048"; do
  echo "== $cfg"; python tools/one_step.py $cfg --ob02 --warm 3 --reps 30 2>&1 | grep -E "wall per build" | cut -c1-200
  python tools/one_step.py $cfg --ob02 --warm 3 --reps 5 --gpu kernel_times=1 2>&1 | grep -E "project_finish" | cut -c1-330
done
