#!/bin/bash
# round 2, call j: persistent lane-refill bisection (ticketed ranges) at 4/5/6 blocks per SM; leftovers in sub-warp groups
set -u
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_ob02.py tests/test_gpu_speculate.py tests/test_gpu_slabs.py -m gpu -x -q > $O/pytest_gpu_j.log 2>&1; echo "pytest exit $?"; tail -3 $O/pytest_gpu_j.log
run() { # lib, cfg..., extra
  local lib=$1; shift
  ISD_LIB_OVERRIDE=$lib python tools/one_step.py "$@" 2>&1 | grep -E "wall per build|project_finish" | cut -c1-330
}
for fs in 0 2; do
  echo "== default lib c4 1024 finish_split=$fs"; run "" c4 1024 --ob02 --warm 3 --reps 30 --gpu finish_split=$fs
  run "" c4 1024 --ob02 --warm 3 --reps 5 --gpu finish_split=$fs kernel_times=1
done
for v in r5 r4; do
  echo "== $v c4 1024 finish_split=2"; run build/variants/lib_$v.so c4 1024 --ob02 --warm 3 --reps 30 --gpu finish_split=2
  run build/variants/lib_$v.so c4 1024 --ob02 --warm 3 --reps 5 --gpu finish_split=2 kernel_times=1
done
for cfg in "c3 512" "c2 256" "c4 2048" "c4 512"; do for fs in 0 2; do
  echo "== $cfg finish_split=$fs"; run "" $cfg --ob02 --warm 3 --reps 20 --gpu finish_split=$fs
done; done
echo "== r5 c3 512 finish_split=2"; run build/variants/lib_r5.so c3 512 --ob02 --warm 3 --reps 20 --gpu finish_split=2
ncu --set full --clock-control none --profile-from-start off -k regex:"k_finish_bisect|k_project_leftovers" -o $O/full_c4_1024_j -f python tools/one_step.py c4 1024 --ob02 --gpu finish_split=2 > $O/ncu_full_c4_j.log 2>&1
ncu -i $O/full_c4_1024_j.ncu-rep --page raw --csv > $O/full_c4_1024_j.csv 2>/dev/null
python tools/ncu_full_summary.py $O/full_c4_1024_j.csv $O/full_c4_1024_j.txt
python tools/ncu_stalls.py $O/full_c4_1024_j.csv 7 | sed 's/smsp__pcsamp_warps_issue_stalled_//g' | cut -c1-400
