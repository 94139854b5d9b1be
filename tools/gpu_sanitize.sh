#!/bin/bash
# compute-sanitizer over small builds of every stage (SURVEY §5 "race detection"): summaries to gpurun_out/sanitize_*.txt
O=gpurun_out; mkdir -p $O
for tool in memcheck racecheck synccheck initcheck; do
  timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_case.py > $O/sanitize_$tool.txt 2>&1
  echo "== $tool exit $?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|built|rank" $O/sanitize_$tool.txt | tail -8
done
