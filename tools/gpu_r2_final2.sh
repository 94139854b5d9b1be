#!/bin/bash
# 2 GPUs: the whole GPU suite (NCCL test included, test_zz last), then bench.py on 2 ranks
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_2gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu_2gpu.log
tail -3 $O/pytest_gpu_2gpu.log
bash tools/gpu_r2_dist.sh 2
