#!/bin/bash
# round 2, call y (2 GPUs): all GPU tests incl. the NCCL test, then bench.py on 2 ranks — halo without the extra layer
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_2gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu_2gpu.log
tail -3 $O/pytest_gpu_2gpu.log
bash tools/gpu_r2_dist.sh 2
