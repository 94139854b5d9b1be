#!/bin/bash
# round 2, call u: units per ticket of the lane-refill kernels (whole mesh and emulated thin slabs)
set -u
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_ob02.py tests/test_gpu_speculate.py -m gpu -x -q 2>&1 | tail -2
for rr in 0 32 64 128; do
  echo "== c4 1024 refill_range=$rr"; python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 30 --gpu refill_range=$rr 2>&1 | grep -E "wall per build" | cut -c1-200
  python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 5 --gpu kernel_times=1 refill_range=$rr 2>&1 | grep -E "eval_tiles=" | cut -c1-330
done
for rr in 32 128; do
  echo "== c3 512 refill_range=$rr"; python tools/one_step.py c3 512 --ob02 --warm 3 --reps 30 --gpu refill_range=$rr 2>&1 | grep -E "wall per build" | cut -c1-200
done
for em in 0/8 3/8; do for rr in 32 128; do
  echo "== emulate $em refill_range=$rr"; python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 3 --emulate $em --gpu kernel_times=1 balance=1 refill_range=$rr 2>&1 | grep -E "eval_tiles=" | cut -c1-330
done; done
