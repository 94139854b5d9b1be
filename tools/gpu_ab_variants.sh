#!/bin/bash
# A/B of library builds (build/variants/lib_*.so) on one headline step: per-kernel times
O=gpurun_out; mkdir -p $O
for v in "$@"; do
  echo "== $v"
  ISD_LIB_OVERRIDE=build/variants/lib_$v.so python tools/one_step.py c4 1024 --ob02 --warm 3 --gpu kernel_times=1 2>&1 | tail -3
done 2>&1 | tee $O/ab_variants.log
