#!/bin/bash
# quick GPU check: parity tests, then timings of the larger configs
set -u
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -15 $O/pytest_gpu.log
timeout 600 python tools/probe.py > $O/probe.log 2>&1; cat $O/probe.log
