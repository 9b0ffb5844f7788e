#!/bin/bash
# usage: tools/gpurun_retry.sh <name> <timeout_s> [gpurun args --] <command...>  — retries while the pod answers "transient"
name=$1; shift; to=$1; shift
log=gpurun_out/call_$name.log
for attempt in 1 2 3 4 5 6 7 8 9 10 11 12; do
  gpurun --timeout $to "$@" > $log 2>&1
  rc=$?
  if grep -q "status=transient" $log || [ $rc -eq 3 ]; then sleep 120; continue; fi
  break
done
echo "done rc=$rc attempts=$attempt" >> $log
