#!/bin/bash
# adjacency from the marching-cubes topology: parity, then A/B
set -u
timeout 600 python -m pytest tests/test_gpu_ob02.py -m gpu -x -q -k "mc_adjacency" 2>&1 | tail -12 | cut -c1-300
for cfg in "c4 1024" "c2 256" "c3 512"; do for t in 1 0; do
  echo "== $cfg mc_adjacency=$t"; python tools/one_step.py $cfg --ob02 --warm 3 --reps 30 --gpu mc_adjacency=$t 2>&1 | grep -E "wall per build" | cut -c1-200
  python tools/one_step.py $cfg --ob02 --warm 3 --reps 5 --gpu kernel_times=1 mc_adjacency=$t 2>&1 | grep -E "eval_tiles=" | cut -c1-420
done; done
