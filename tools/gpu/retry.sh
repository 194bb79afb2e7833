#!/bin/bash
# usage: tools/gpu/retry.sh <timeout> <logfile> [gpurun options, e.g. --gpus 2] -- <command>
# retries gpurun while the pod answers busy (transient / no box or slot free)
T=$1; LOG=$2; shift 2
OPTS=()
while [ "$1" != "--" ] && [ $# -gt 0 ]; do OPTS+=("$1"); shift; done
shift
for i in $(seq 1 14); do
  /usr/local/graft/bin/gpurun --timeout $T "${OPTS[@]}" -- "$@" > $LOG 2>&1
  if grep -q "status=transient\|no box or slot\|status=busy" $LOG; then sleep 100; continue; fi
  break
done
tail -40 $LOG
