#!/bin/bash
# closing GPU call of a session: parity tests, smoke, bench, stage times of the large configs, launch list and full ncu
# capture of one C2 step, full capture of the OB02 kernels of one C4 1024^3 step (reps converted to CSV on the box)
set -u
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 300 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; cat $O/bench.json
rm -f $O/proj_times.log
for cfg in "c2 256" "c4 512" "c4 1024" "c3 512" "c4 2048"; do
  timeout 200 python tools/one_step.py $cfg --ob02 --warm 3 --reps 1 2>&1 | tail -1 | tee -a $O/proj_times.log
done
python tools/one_step.py c2 256 --ob02 > $O/plain_c2.log 2>&1 &&
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $O/launches_c2_256.csv python tools/one_step.py c2 256 --ob02 > $O/ncu_c2.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on --profile-from-start off -o $O/full_c2_256 -f python tools/one_step.py c2 256 --ob02 > $O/ncu_full_c2.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'k_build_point_lists|k_centroid_gradients|k_project|k_qem|k_face|k_eval_tiles|k_mc_emit' -o $O/full_c4_1024 -f python tools/one_step.py c4 1024 --ob02 > $O/ncu_full_c4.log 2>&1
for t in c2_256 c4_1024; do ncu -i $O/full_$t.ncu-rep --page raw --csv > $O/full_$t.csv 2>/dev/null; rm -f $O/full_$t.ncu-rep; done
ls -la $O | head -30
