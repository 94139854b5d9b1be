#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_ob02.py tests/test_gpu_speculate.py tests/test_gpu_slabs.py tests/test_gpu_sizes.py -m gpu -x -q 2>&1 | tail -2
python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 30 2>&1 | grep -E "wall per build"
