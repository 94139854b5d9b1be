#!/bin/bash
# round 2, call b: parity tests with the packed interpreters, bench without CPU legs, ncu of one headline step
set -u
O=gpurun_out
mkdir -p $O
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp32_peak tools/fp32_peak.cu 2>/dev/null; tools/fp32_peak > $O/fp32_peak.json; cat $O/fp32_peak.json
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -15 $O/pytest_gpu.log
timeout 900 python bench.py --no-cpu-baseline > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print(d['ms_per_step'], d['value'], d['e2e']['value'], d['stage_ms'])
for k,v in d['kernels'].items(): print('%-70s %.4f ms  %s %s'%(k, v['ms_per_step'], v.get('achieved'), v.get('unit')))
print({k:(v.get('ms_per_step')) for k,v in d['other_configs'].items()})
PY
tail -5 $O/bench.err
for a in "pair=0" "pair=1"; do python tools/one_step.py c4 1024 --ob02 --gpu $a; done > $O/ab_pair.log 2>&1; cat $O/ab_pair.log
python tools/one_step.py c4 1024 --ob02 > $O/plain_c4.log 2>&1 &&
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'k_eval_tiles|pair' -o $O/full_c4_1024_b -f python tools/one_step.py c4 1024 --ob02 > $O/ncu_full_c4.log 2>&1
ncu -i $O/full_c4_1024_b.ncu-rep --page raw --csv > $O/full_c4_1024_b.csv 2>/dev/null
ls -la $O
