#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -6 $O/pytest_gpu.log
bash tools/gpu_ab_variants.sh default mb4 mb5 mb6 mb5_noinline n8 n8mb8
echo "== default, finish_split=0"; python tools/one_step.py c4 1024 --ob02 --warm 3 --gpu kernel_times=1 finish_split=0 2>&1 | tail -2
