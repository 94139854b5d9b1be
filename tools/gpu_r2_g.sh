#!/bin/bash
# round 2, call g: GPU tests, full bench (both arms), compute-sanitizer summaries
set -u
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
timeout 900 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench exit $?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref exit $?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print(d['ms_per_step'], d['value'], d['e2e']['value'], d['stage_ms'])
for k,v in d['kernels'].items(): print('%-70s %.4f ms  %s %s'%(k, v['ms_per_step'], v.get('achieved'), v.get('unit')))
print({k:(v.get('ms_per_step')) for k,v in d['other_configs'].items()})
print(d.get('cpu_baseline'))
PY
for tool in memcheck racecheck synccheck initcheck; do
  timeout 600 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_case.py > $O/sanitize_$tool.txt 2>&1
  echo "== $tool exit $?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|built|rank" $O/sanitize_$tool.txt | tail -8
done
