#!/bin/bash
# parity tests + stage times (optionally against another build: ISD_AB_OLD=path/to/lib.so)
set -u
O=gpurun_out; mkdir -p $O; rm -f $O/ab.log
timeout 200 python -m pytest tests -m gpu -x -q 2>&1 | tail -2 | tee -a $O/ab.log
for cfg in ${ISD_AB_CFGS:-"c2 256" "c4 1024" "c4 2048"}; do :; done
for cfg in "c2 256" "c4 1024" "c3 512"; do
  timeout 200 python tools/one_step.py $cfg --ob02 --warm 3 --reps 1 2>&1 | tail -1 | tee -a $O/ab.log
done
