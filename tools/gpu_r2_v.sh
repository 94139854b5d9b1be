#!/bin/bash
# round 2, call v: why is the trials kernel slow on the upper half slab? (emulated rank 1/2)
set -u
O=gpurun_out; mkdir -p $O
for extra in "refill_range=64" "refill_range=32" "search_split=0"; do
  echo "== emulate 1/2 $extra"; python tools/one_step.py c4 1024 --ob02 --warm 3 --reps 3 --emulate 1/2 --gpu kernel_times=1 balance=1 $extra 2>&1 | grep -E "eval_tiles=|rror" | cut -c1-330
done
ncu --set full --clock-control none --profile-from-start off -k regex:"k_search_trials" -o $O/full_em12_v -f python tools/one_step.py c4 1024 --ob02 --emulate 1/2 --gpu balance=1 > $O/ncu_v.log 2>&1
ncu -i $O/full_em12_v.ncu-rep --page raw --csv > $O/full_em12_v.csv 2>/dev/null
python tools/ncu_full_summary.py $O/full_em12_v.csv $O/full_em12_v.txt | cut -c1-150
python tools/ncu_stalls.py $O/full_em12_v.csv 7 | sed 's/smsp__pcsamp_warps_issue_stalled_//g' | cut -c1-400
