#!/bin/bash
# usage: tools/grun.sh <log name> <timeout s> <command...>   — gpurun with retries while the pod answers "busy" (exit 3)
log=gpurun_out/$1.log; to=$2; shift 2
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $to -- "$@" > $log 2>&1; rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" $log; then break; fi
  sleep 45
done
echo "grun done rc=$rc" >> $log
