#!/bin/bash
# usage: tools/gpu_retry.sh <logfile> <timeout> <command...>   -- retries gpurun while the pod answers "busy" (nothing charged)
LOG=$1; TO=$2; shift 2
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun --timeout $TO -- "$@" > $LOG 2>&1
  if ! grep -q "status=transient" $LOG; then exit 0; fi
  sleep 90
done
