#!/bin/bash
# round 2, call h: adjacency on a side stream (A/B), cost of kernel_times, gaps, full ncu of the headline step with stall reasons
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_ob02.py tests/test_gpu_speculate.py tests/test_gpu_slabs.py tests/test_gpu_next_rows.py -m gpu -x -q > $O/pytest_gpu_h.log 2>&1; echo "pytest exit $?"; tail -3 $O/pytest_gpu_h.log
for extra in "overlap=1" "overlap=0" "overlap=1 kernel_times=1" "overlap=0 kernel_times=1"; do
  echo "== c4 1024 $extra"; python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 30 --gpu $extra 2>&1 | tail -4 | cut -c1-600
done
for cfg in "c2 256" "c3 512"; do for extra in "overlap=1" "overlap=0"; do
  echo "== $cfg $extra"; python tools/one_step.py $cfg --ob02 --warm 3 --reps 30 --gpu $extra 2>&1 | tail -2 | cut -c1-400
done; done
python tools/one_step.py c4 1024 --ob02 > $O/plain_c4.log 2>&1 &&
ncu --set full --clock-control none --import-source on --profile-from-start off -o $O/full_c4_1024_h -f python tools/one_step.py c4 1024 --ob02 --gpu overlap=0 > $O/ncu_full_c4.log 2>&1
ncu -i $O/full_c4_1024_h.ncu-rep --page raw --csv > $O/full_c4_1024_h.csv 2>/dev/null
tail -2 $O/ncu_full_c4.log
python tools/ncu_full_summary.py $O/full_c4_1024_h.csv $O/full_c4_1024_h.txt > /dev/null
python tools/ncu_stalls.py $O/full_c4_1024_h.csv 7 > $O/stalls_c4_1024_h.txt; head -c 6000 $O/stalls_c4_1024_h.txt
