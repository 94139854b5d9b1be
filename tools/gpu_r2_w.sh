#!/bin/bash
# round 2, call w: ghost-layer vertices written in halo mode
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
for em in 1/2 3/8 0/8; do
  echo "== emulate $em"; python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 3 --emulate $em --gpu kernel_times=1 balance=1 2>&1 | grep -E "eval_tiles=|rror|^c4" | cut -c1-400
done
echo "== whole"; python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 30 2>&1 | grep -E "wall per build" | cut -c1-200
