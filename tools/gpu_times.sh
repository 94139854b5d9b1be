#!/bin/bash
# per-launch device times of one step: gpu_times.sh <tag> <one_step args...>
set -u
O=gpurun_out; mkdir -p $O
tag=$1; shift
python tools/one_step.py "$@" > $O/plain_$tag.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $O/launches_$tag.csv python tools/one_step.py "$@" > $O/ncu_$tag.log 2>&1
python - <<PY
import csv
rows=list(csv.DictReader(l for l in open('$O/launches_$tag.csv') if not l.startswith('==')))
tot=0
for r in rows:
    n=r['Kernel Name'].split('(')[0].replace('isd::','').replace('void ','')
    v=float(r['Metric Value'].replace(',',''))/1000; tot+=v
    print('%-36s %9.1f us'%(n[:36],v))
print('total %.1f us'%tot)
PY
