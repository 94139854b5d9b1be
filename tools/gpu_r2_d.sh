#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_ob02.py -m gpu -x -q > $O/pytest_gpu_ob02.log 2>&1; echo "pytest exit $?"; tail -3 $O/pytest_gpu_ob02.log
bash tools/gpu_ab_variants.sh default n8mb8_p6 n8mb10_p7 n8mb12_p8 n4mb8_p6n
echo "== n8mb8_p6, finish_split=0"; ISD_LIB_OVERRIDE=build/variants/lib_n8mb8_p6.so python tools/one_step.py c4 1024 --ob02 --warm 3 --gpu kernel_times=1 finish_split=0 2>&1 | tail -3
