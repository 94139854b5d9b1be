#!/bin/bash
# round 2, call m: faces-of-faces one thread per vertex
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
for cfg in "c4 1024" "c2 256"; do for fof in 1 0; do
  echo "== $cfg fof=$fof"; python tools/one_step.py $cfg --ob02 --warm 3 --reps 30 --gpu fof=$fof 2>&1 | grep -E "wall per build" | cut -c1-330
  python tools/one_step.py $cfg --ob02 --warm 3 --reps 5 --gpu kernel_times=1 fof=$fof 2>&1 | grep -E "project_finish" | cut -c1-330
done; done
