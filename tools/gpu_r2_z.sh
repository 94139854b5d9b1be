#!/bin/bash
# the no-allocation entry point: its test alone in a fresh process, then the whole GPU suite (it is collected last)
set -u
O=gpurun_out; mkdir -p $O
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/pytest_noalloc_alone.log 2>&1; echo "alone exit $?"; tail -4 $O/pytest_noalloc_alone.log | cut -c1-300
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log | cut -c1-300
