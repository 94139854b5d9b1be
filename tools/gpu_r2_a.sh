#!/bin/bash
# round 2, first GPU call: parity tests, bench (headline C4 1024^3), reference arm, launch list + full ncu capture of one headline step
set -u
O=gpurun_out
mkdir -p $O
nproc > $O/host.txt; free -g | head -2 >> $O/host.txt; nvidia-smi -L >> $O/host.txt
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp32_peak tools/fp32_peak.cu 2>/dev/null
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -5 $O/pytest_gpu.log
timeout 900 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; cat $O/bench.json; tail -5 $O/bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2>> $O/bench.err; cat $O/bench_ref.json
python tools/one_step.py c4 1024 --ob02 > $O/plain_c4.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $O/launches_c4_1024.csv python tools/one_step.py c4 1024 --ob02 > $O/ncu_c4.log 2>&1
python tools/one_step.py c4 1024 --ob02 > $O/plain_c4.log 2>&1 &&
ncu --set full --clock-control none --import-source on --profile-from-start off -o $O/full_c4_1024 -f python tools/one_step.py c4 1024 --ob02 > $O/ncu_full_c4.log 2>&1
ncu -i $O/full_c4_1024.ncu-rep --page raw --csv > $O/full_c4_1024.csv 2>/dev/null
cat $O/plain_c4.log
ls -la $O
