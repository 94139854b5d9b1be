#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_speculate.py tests/test_gpu_ob02.py -m gpu -x -q > $O/pytest_gpu_e.log 2>&1; echo "pytest exit $?"; tail -5 $O/pytest_gpu_e.log
bash tools/gpu_ab_variants.sh n8mb8_p6 n8mb8_p5 n8mb8_p7 n8mb8_p6n n4mb8_p6 n8mb10_p6
for extra in "finish_split=0" "speculate=0" "pair=0"; do
echo "== n8mb8_p6, $extra"; ISD_LIB_OVERRIDE=build/variants/lib_n8mb8_p6.so python tools/one_step.py c4 1024 --ob02 --warm 3 --gpu kernel_times=1 $extra 2>&1 | tail -3
done
echo "== n8mb8_p6 host wall, no kernel_times"; ISD_LIB_OVERRIDE=build/variants/lib_n8mb8_p6.so python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 20 2>&1 | tail -1
echo "== n8mb8_p6 host wall, no kernel_times, speculate=0"; ISD_LIB_OVERRIDE=build/variants/lib_n8mb8_p6.so python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 20 --gpu speculate=0 2>&1 | tail -1
