#!/bin/bash
# round 2, call p: quad centroid gradients A/B; where does the thin-slab cost come from (flops per face in emulated slabs)
set -u
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_ob02.py -m gpu -x -q 2>&1 | tail -2
for lib in "" build/variants/lib_q5.so; do for q in 0 1; do
  echo "== lib=$lib quad=$q"; ISD_LIB_OVERRIDE=$lib python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 30 --gpu quad=$q 2>&1 | grep -E "wall per build" | cut -c1-200
  ISD_LIB_OVERRIDE=$lib python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 5 --gpu quad=$q kernel_times=1 2>&1 | grep -E "project_finish" | cut -c1-200
done; done
echo "== whole"; python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 3 --flops --gpu kernel_times=1 2>&1 | grep -E "^flops|project_finish=|nodes" | cut -c1-900
for em in 0/8 3/8 4/8 7/8; do
  echo "== emulate $em (static cuts)"; python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 3 --emulate $em --flops --gpu kernel_times=1 balance=1 2>&1 | grep -E "^flops|project_finish=|nodes" | cut -c1-900
done
