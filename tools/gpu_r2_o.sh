#!/bin/bash
# round 2, call o: list-entry prefetch A/B + parity; emulated half-grid throughput
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
for lib in "" build/variants/lib_nopf.so; do
  echo "== lib=$lib c4 1024"; ISD_LIB_OVERRIDE=$lib python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 30 2>&1 | grep -E "wall per build" | cut -c1-330
  ISD_LIB_OVERRIDE=$lib python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 5 --gpu kernel_times=1 2>&1 | grep -E "project_finish" | cut -c1-330
  echo "== lib=$lib c3 512"; ISD_LIB_OVERRIDE=$lib python tools/one_step.py c3 512 --ob02 --warm 3 --reps 30 2>&1 | grep -E "wall per build" | cut -c1-330
done
for em in 0/2 1/2 0/8 3/8; do
  echo "== emulate $em"; python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 5 --emulate $em --gpu kernel_times=1 2>&1 | grep -E "project_finish|nodes" | cut -c1-420
done
