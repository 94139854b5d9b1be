#!/bin/bash
# round 2: GPU tests, bench (both arms), ncu launch list and full capture of one headline step
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; tail -3 $O/bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref exit $?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print(d['ms_per_step'], d['value'], d['e2e']['value'], d['instrumented']['ms_per_step'], d['stage_ms'])
for k,v in d['kernels'].items(): print('%-70s %.4f ms  %s %s frac %s'%(k, v['ms_per_step'], v.get('achieved'), v.get('unit'), v.get('frac')))
print({k:(v.get('ms_per_step')) for k,v in d['other_configs'].items()})
print({k:v for k,v in d['cpu_baseline'].items() if k!='sample'})
PY
python tools/one_step.py c4 1024 --ob02 > $O/plain_c4.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $O/launches_c4_1024.csv python tools/one_step.py c4 1024 --ob02 > $O/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on --profile-from-start off -o $O/full_c4_1024_final -f python tools/one_step.py c4 1024 --ob02 > $O/ncu_full_c4.log 2>&1
ncu -i $O/full_c4_1024_final.ncu-rep --page raw --csv > $O/full_c4_1024_final.csv 2>/dev/null
python tools/ncu_full_summary.py $O/full_c4_1024_final.csv $O/full_c4_1024_final.txt | cut -c1-150
python tools/ncu_stalls.py $O/full_c4_1024_final.csv 6 | sed 's/smsp__pcsamp_warps_issue_stalled_//g' | cut -c1-330 > $O/stalls_c4_1024_final.txt
