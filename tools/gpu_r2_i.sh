#!/bin/bash
# round 2, call i: lane-refill bisection (finish_split=2) against the one-kernel finish and the batch form
set -u
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_ob02.py tests/test_gpu_speculate.py -m gpu -x -q > $O/pytest_gpu_i.log 2>&1; echo "pytest exit $?"; tail -3 $O/pytest_gpu_i.log
for extra in "finish_split=0" "finish_split=1" "finish_split=2"; do
  echo "== c4 1024 $extra"; python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 30 --gpu $extra 2>&1 | tail -2 | cut -c1-400
  echo "== c4 1024 $extra kernel_times=1"; python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 5 --gpu $extra kernel_times=1 2>&1 | tail -3 | head -1 | cut -c1-400
done
for cfg in "c3 512" "c4 2048"; do for extra in "finish_split=0" "finish_split=2"; do
  echo "== $cfg $extra"; python tools/one_step.py $cfg --ob02 --warm 3 --reps 10 --gpu $extra 2>&1 | tail -2 | cut -c1-400
done; done
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"k_finish_bisect|k_unit_gradients" -o $O/full_c4_1024_i -f python tools/one_step.py c4 1024 --ob02 --gpu finish_split=2 > $O/ncu_full_c4_i.log 2>&1
ncu -i $O/full_c4_1024_i.ncu-rep --page raw --csv > $O/full_c4_1024_i.csv 2>/dev/null
python tools/ncu_full_summary.py $O/full_c4_1024_i.csv $O/full_c4_1024_i.txt
python tools/ncu_stalls.py $O/full_c4_1024_i.csv 7 | sed 's/smsp__pcsamp_warps_issue_stalled_//g' | cut -c1-400
ncu --set full --clock-control none --profile-from-start off -k regex:"k_finish_bisect|k_unit_gradients" -o $O/full_c4_1024_i1 -f python tools/one_step.py c4 1024 --ob02 --gpu finish_split=1 > $O/ncu_full_c4_i1.log 2>&1
ncu -i $O/full_c4_1024_i1.ncu-rep --page raw --csv > $O/full_c4_1024_i1.csv 2>/dev/null
python tools/ncu_full_summary.py $O/full_c4_1024_i1.csv $O/full_c4_1024_i1.txt
python tools/ncu_stalls.py $O/full_c4_1024_i1.csv 7 | sed 's/smsp__pcsamp_warps_issue_stalled_//g' | cut -c1-400
