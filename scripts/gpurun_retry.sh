#!/bin/bash
# usage: scripts/gpurun_retry.sh <timeout> <logfile> <command...>  - retries while the pod answers "transient"
to=$1; log=$2; shift 2
for i in $(seq 1 30); do
  gpurun --timeout $to -- "$@" > $log 2>&1
  if grep -q "status=transient\|status=busy" $log; then sleep 90; else break; fi
done
