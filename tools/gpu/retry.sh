#!/bin/bash
# usage: tools/gpu/retry.sh <timeout> <logfile> <command...>   -- retries gpurun while the pod answers busy (rc 3 / transient)
T=$1; LOG=$2; shift 2
for i in $(seq 1 12); do
  /usr/local/graft/bin/gpurun --timeout $T -- "$@" > $LOG 2>&1
  if grep -q "status=transient\|no box or slot" $LOG; then sleep 100; continue; fi
  break
done
tail -40 $LOG
