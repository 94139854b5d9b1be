#!/bin/bash
# round 2, call x: occupancy of the gradient-only pair kernels (blocks per SM)
set -u
for v in "" g8 rf7 cg8; do
  lib=""; [ -n "$v" ] && lib=build/variants/lib_$v.so
  echo "== variant '$v' c4 1024"; ISD_LIB_OVERRIDE=$lib python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 30 2>&1 | grep -E "wall per build" | cut -c1-200
  ISD_LIB_OVERRIDE=$lib python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 5 --gpu kernel_times=1 2>&1 | grep -E "eval_tiles=" | cut -c1-260
  echo "== variant '$v' c3 512"; ISD_LIB_OVERRIDE=$lib python tools/one_step.py c3 512 --ob02 --warm 3 --reps 30 2>&1 | grep -E "wall per build" | cut -c1-200
done
