#!/bin/bash
# round 2, call k: all GPU tests + bench at the new defaults (lane-refill finish, grouped leftovers)
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
timeout 900 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; tail -3 $O/bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print(d['ms_per_step'], d['value'], d['e2e']['value'], d['instrumented']['ms_per_step'], d['stage_ms'])
for k,v in d['kernels'].items(): print('%-70s %.4f ms  %s %s frac %s'%(k, v['ms_per_step'], v.get('achieved'), v.get('unit'), v.get('frac')))
print({k:(v.get('ms_per_step')) for k,v in d['other_configs'].items()})
print({k:v for k,v in d['cpu_baseline'].items() if k!='sample'})
PY
