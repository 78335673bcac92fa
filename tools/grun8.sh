#!/bin/bash
# usage: tools/grun8.sh <log name> <timeout s> <command...>   — 8-GPU gpurun with retries while the pod answers "busy"
log=gpurun_out/$1.log; to=$2; shift 2
for i in $(seq 1 60); do
  /usr/local/graft/bin/gpurun --gpus 8 --timeout $to -- "$@" > $log 2>&1; rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" $log; then break; fi
  sleep 60
done
echo "grun done rc=$rc" >> $log
