#!/bin/bash
set -u
for ps in -1 1; do
echo "== c3ii 512 point_spec=$ps"; python tools/one_step.py c3ii 512 --ob02 --warm 3 --reps 20 --gpu point_spec=$ps 2>&1 | grep -E "wall per build" | cut -c1-330
python tools/one_step.py c3ii 512 --ob02 --warm 3 --reps 5 --gpu kernel_times=1 point_spec=$ps 2>&1 | grep -E "eval_tiles=" | cut -c1-400
done
