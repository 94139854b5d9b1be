#!/bin/bash
# round 2, call f: all GPU tests with the baked defaults + speculative builds, bench (N=1, no CPU legs), full ncu of one headline step
set -u
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -8 $O/pytest_gpu.log
timeout 900 python bench.py --no-cpu-baseline > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print(d['ms_per_step'], d['value'], d['e2e']['value'], d['stage_ms'], d['speculation'])
for k,v in d['kernels'].items(): print('%-70s %.4f ms  %s %s'%(k, v['ms_per_step'], v.get('achieved'), v.get('unit')))
print({k:(v.get('ms_per_step')) for k,v in d['other_configs'].items()})
PY
tail -5 $O/bench.err
python tools/one_step.py c4 1024 --ob02 > $O/plain_c4.log 2>&1 &&
ncu --set full --clock-control none --import-source on --profile-from-start off -o $O/full_c4_1024_f -f python tools/one_step.py c4 1024 --ob02 > $O/ncu_full_c4.log 2>&1
ncu -i $O/full_c4_1024_f.ncu-rep --page raw --csv > $O/full_c4_1024_f.csv 2>/dev/null
tail -3 $O/ncu_full_c4.log
