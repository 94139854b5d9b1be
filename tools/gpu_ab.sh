#!/bin/bash
# parity tests + stage times of a few configs; ISD_AB_OLD=path/to/another/libimplisolid_b200.so adds the same timings for that build
set -u
O=gpurun_out; mkdir -p $O; rm -f $O/ab.log
timeout 200 python -m pytest tests -m gpu -x -q 2>&1 | tail -2 | tee -a $O/ab.log
for cfg in "c2 256" "c4 1024" "c3 512"; do
  timeout 200 python tools/one_step.py $cfg --ob02 --warm 3 --reps 1 2>&1 | tail -1 | tee -a $O/ab.log
  if [ -n "${ISD_AB_OLD:-}" ]; then
    ISD_LIB_OVERRIDE=$ISD_AB_OLD timeout 200 python tools/one_step.py $cfg --ob02 --warm 3 --reps 1 2>&1 | tail -1 | sed 's/^/OLD /' | tee -a $O/ab.log
  fi
done
