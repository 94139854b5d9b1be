#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
N=${1:-1}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29811 --nproc-per-node $N"
timeout 300 $TR tools/dist_scale.py c4 2048 --ob02 --steps 4 2>> $O/scale.err | grep '^{' | tee -a $O/scale_c5.jsonl
if [ "$N" != "1" ]; then timeout 300 $TR tools/dist_scale.py c4 1024 --ob02 --steps 6 2>> $O/scale.err | grep '^{' | tee -a $O/scale_c5.jsonl; timeout 300 $TR bench.py --gpus $N --steps 10 --warmup 3 2>> $O/scale.err | grep '^{' | tee -a $O/scale_c5.jsonl; else timeout 300 python tools/dist_scale.py c4 1024 --ob02 --steps 6 | grep '^{' | tee -a $O/scale_c5.jsonl; fi
