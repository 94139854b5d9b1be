#!/bin/bash
# usage: gpu_r2_dist.sh N   : NCCL parity check on N ranks, then bench.py at N ranks (headline C4 1024^3 + c5_2048 record)
set -u
N=$1
O=gpurun_out
mkdir -p $O
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 tools/dist_check.py > $O/dist_check_$N.log 2>&1; echo "dist_check exit $?"
grep "^{" $O/dist_check_$N.log | tail -1 | cut -c1-3000; tail -3 $O/dist_check_$N.log | cut -c1-600
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus $N --steps 50 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench exit $?"
tail -3 $O/bench_n$N.err | cut -c1-600
python - <<PY
import json
for l in open('$O/bench_n$N.json'):
    if l.startswith('{'):
        d=json.loads(l)
        print('N', d['n_gpus'], 'ms', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'], 'parity', d.get('parity'))
        print('stage max', d.get('stage_ms_max_over_ranks')); print('stage min', d.get('stage_ms_min_over_ranks')); print('limiting', d.get('limiting_stage'))
        for r in d.get('per_rank', []): print(r['rank'], r['z'], r['nv'], r['nf'], {k: round(v,3) for k,v in r['stage_ms'].items() if v})
        print({k:(v.get('ms_per_step'), v.get('value')) for k,v in d['other_configs'].items()})
PY
if [ "${AB_FINISH:-0}" = "1" ]; then
for extra in "finish_split=0" "finish_split=1"; do echo "== $extra"; python tools/one_step.py c4 1024 --ob02 --warm 3 --gpu kernel_times=1 $extra 2>&1 | tail -3 | cut -c1-400; done
timeout 600 python -m pytest tests/test_gpu_ob02.py -m gpu -x -q -k "pair_kernels" 2>&1 | tail -3
fi
