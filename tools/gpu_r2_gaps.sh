#!/bin/bash
N=$1; O=gpurun_out; mkdir -p $O
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29621 tools/dist_gaps.py c4 1024 > $O/dist_gaps_$N.log 2>&1; echo "exit $?"
grep -E "^rank|^   (kernels|gaps)" $O/dist_gaps_$N.log | cut -c1-420
