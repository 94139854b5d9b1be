#!/bin/bash
set -u
for v in "" emitnocap; do
  lib=""; [ -n "$v" ] && lib=build/variants/lib_$v.so
  echo "== variant '$v' c4 1024"; ISD_LIB_OVERRIDE=$lib python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 30 2>&1 | grep -E "wall per build" | cut -c1-200
  ISD_LIB_OVERRIDE=$lib python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 5 --gpu kernel_times=1 2>&1 | grep -E "eval_tiles=" | tr ' ' '\n' | grep -E "adjacency|mc_emit" | tr '\n' ' '; echo
done
timeout 300 python -m pytest tests/test_gpu_ob02.py -m gpu -x -q -k "mc_adjacency" 2>&1 | tail -1
