#!/bin/bash
# first GPU call of a session: parity tests, bench, launch list, full ncu capture of one step, config sweep
set -u
O=gpurun_out
mkdir -p $O
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp32_peak tools/fp32_peak.cu 2>/dev/null
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; cat $O/bench.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2>> $O/bench.err; cat $O/bench_ref.json
timeout 600 python tools/probe.py > $O/probe.log 2>&1; cat $O/probe.log
python tools/one_step.py c2 256 --ob02 > $O/plain_c2.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $O/launches_c2_256.csv python tools/one_step.py c2 256 --ob02 > $O/ncu_c2.log 2>&1
python tools/one_step.py c2 256 --ob02 > $O/plain_c2.log 2>&1 &&
ncu --set full --clock-control none --import-source on --profile-from-start off -o $O/full_c2_256 -f python tools/one_step.py c2 256 --ob02 > $O/ncu_full_c2.log 2>&1
python tools/one_step.py c4 1024 --ob02 > $O/plain_c4.log 2>&1 &&
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'k_super_bounds|k_tile_bounds|k_eval_tiles|k_mc_classify|k_row_scan|k_seg_apply|k_mc_emit|k_centroid_gradients|k_project|k_qem|k_face' -o $O/full_c4_1024 -f python tools/one_step.py c4 1024 --ob02 > $O/ncu_full_c4.log 2>&1
for t in c2_256 c4_1024; do ncu -i $O/full_$t.ncu-rep --page raw --csv > $O/full_$t.csv 2>/dev/null; rm -f $O/full_$t.ncu-rep; done
cat $O/plain_c4.log
ls -la $O
