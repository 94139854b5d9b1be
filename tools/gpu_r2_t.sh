#!/bin/bash
# round 2, call t: the search's trial points in a lane-refill kernel
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
for cfg in "c4 1024" "c3 512" "c4 2048"; do for ss in 1 0; do
  echo "== $cfg search_split=$ss"; python tools/one_step.py $cfg --ob02 --warm 3 --reps 30 --gpu search_split=$ss 2>&1 | grep -E "wall per build" | cut -c1-300
  python tools/one_step.py $cfg --ob02 --warm 3 --reps 5 --gpu kernel_times=1 search_split=$ss 2>&1 | grep -E "eval_tiles=" | cut -c1-420
done; done
for v in t5 r5 t5r5; do
  echo "== variant $v c4 1024"; ISD_LIB_OVERRIDE=build/variants/lib_$v.so python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 30 2>&1 | grep -E "wall per build" | cut -c1-300
  ISD_LIB_OVERRIDE=build/variants/lib_$v.so python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 5 --gpu kernel_times=1 2>&1 | grep -E "eval_tiles=" | cut -c1-300
  echo "== variant $v c3 512"; ISD_LIB_OVERRIDE=build/variants/lib_$v.so python tools/one_step.py c3 512 --ob02 --warm 3 --reps 30 2>&1 | grep -E "wall per build" | cut -c1-300
done
ncu --set full --clock-control none --profile-from-start off -k regex:"k_search_trials|k_project_search" -o $O/full_c4_1024_t -f python tools/one_step.py c4 1024 --ob02 > $O/ncu_full_c4_t.log 2>&1
ncu -i $O/full_c4_1024_t.ncu-rep --page raw --csv > $O/full_c4_1024_t.csv 2>/dev/null
python tools/ncu_full_summary.py $O/full_c4_1024_t.csv $O/full_c4_1024_t.txt | cut -c1-150
python tools/ncu_stalls.py $O/full_c4_1024_t.csv 5 | sed 's/smsp__pcsamp_warps_issue_stalled_//g' | cut -c1-330
