#!/bin/bash
# last validation of the round: GPU suite, smoke, bench (own arm)
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; tail -3 $O/bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print(d['ms_per_step'], d['value'], d['e2e']['value'], d['instrumented']['ms_per_step'])
for k,v in list(d['kernels'].items())[:8]: print('%-70s %.4f ms  %s %s frac %s'%(k, v['ms_per_step'], v.get('achieved'), v.get('unit'), v.get('frac')))
print({k:(v.get('ms_per_step')) for k,v in d['other_configs'].items()})
print({k:v for k,v in d['cpu_baseline'].items() if k not in ('sample','stage_s')})
PY
