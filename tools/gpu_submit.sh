#!/bin/bash
# usage: tools/gpu_submit.sh <log> <timeout_s> [--gpus N] -- <command>   : retries while the pod answers "busy" (exit 3)
log=$1; shift; to=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $to "$@" > $log 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then break; fi
  sleep 90
done
echo "submit rc=$rc tries=$i" >> $log
