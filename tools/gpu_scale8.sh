#!/bin/bash
# 8-GPU box: NCCL parity, z-slab scaling of C4 1024^3 (eval+MC) and C5 2048^3 (eval+MC+OB02), bench.py at N=2,4,8
set -u
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
: > $O/scale.jsonl
port=29700
run() { port=$((port+1)); n=$1; shift; timeout 300 $TR --master-port $port --nproc-per-node $n "$@" 2>> $O/scale.err | grep '^{' | tee -a $O/scale.jsonl; }
timeout 600 python -m pytest tests/test_gpu_nccl.py -x -q 2>&1 | tail -3
for n in ${SCALE_C4:-8}; do run $n tools/dist_scale.py c4 1024 --steps 6; done
for n in ${SCALE_C5:-2 4 8}; do run $n tools/dist_scale.py c4 2048 --ob02 --steps 4; done
for n in ${SCALE_BENCH:-2 4 8}; do run $n bench.py --gpus $n --steps 10 --warmup 3; done
grep -v "^\*\*\*\|OMP_NUM\|^$" $O/scale.err | tail -5
