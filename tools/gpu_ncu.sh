#!/bin/bash
# usage: gpu_ncu.sh <tag> <kernel regex> <one_step args...>   -> gpurun_out/full_<tag>.ncu-rep
set -u
O=gpurun_out; mkdir -p $O
tag=$1; shift; rx=$1; shift
python tools/one_step.py "$@" > $O/plain_$tag.log 2>&1 &&
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"$rx" -o $O/full_$tag -f python tools/one_step.py "$@" > $O/ncu_$tag.log 2>&1
cat $O/plain_$tag.log; tail -2 $O/ncu_$tag.log
