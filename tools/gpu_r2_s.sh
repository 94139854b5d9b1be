#!/bin/bash
# round 2, call s: faces-of-faces by umbrella intersection; bisection step counters
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
for cfg in "c4 1024" "c3 512" "c2 256" "c4 2048"; do
  echo "== $cfg"; python tools/one_step.py $cfg --ob02 --warm 3 --reps 30 2>&1 | grep -E "wall per build|^bisected" | cut -c1-300
  python tools/one_step.py $cfg --ob02 --warm 3 --reps 5 --gpu kernel_times=1 2>&1 | grep -E "eval_tiles=" | cut -c1-420
done
