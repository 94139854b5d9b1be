#!/bin/bash
# round 2, call r: faces grouped by tile for the pair kernels (perm)
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
for cfg in "c4 1024" "c3 512" "c2 256" "c4 2048"; do for perm in 1 0; do
  echo "== $cfg perm=$perm"; python tools/one_step.py $cfg --ob02 --warm 3 --reps 30 --gpu perm=$perm 2>&1 | grep -E "wall per build" | cut -c1-200
  python tools/one_step.py $cfg --ob02 --warm 3 --reps 5 --gpu kernel_times=1 perm=$perm 2>&1 | grep -E "eval_tiles=" | cut -c1-380
done; done
ncu --set full --clock-control none --profile-from-start off -k regex:"k_finish_bisect|k_unit_gradients|k_project_search|k_centroid_gradients|k_seg_apply" -o $O/full_c4_1024_r -f python tools/one_step.py c4 1024 --ob02 > $O/ncu_full_c4_r.log 2>&1
ncu -i $O/full_c4_1024_r.ncu-rep --page raw --csv > $O/full_c4_1024_r.csv 2>/dev/null
python tools/ncu_full_summary.py $O/full_c4_1024_r.csv $O/full_c4_1024_r.txt | cut -c1-150
python tools/ncu_stalls.py $O/full_c4_1024_r.csv 5 | sed 's/smsp__pcsamp_warps_issue_stalled_//g' | cut -c1-330
