#!/bin/bash
mkdir -p gpurun_out
run() { # label env...
  local label=$1; shift
  for g in 1 2 4 8; do
    env "$@" timeout 300 python tools/rank_emul.py --gpus $g --frames 5 > gpurun_out/r2o_${label}_n${g}.log 2>&1
    echo "$label n$g $(tail -n 1 gpurun_out/r2o_${label}_n${g}.log | cut -c16-70)"
  done
}
run default RTC_X=0
run serialflag RTC_PATHS=serial
run poollast RTC_POOL_LAST=1
run carve RTC_CARVEOUT=1
run both RTC_POOL_LAST=1 RTC_CARVEOUT=1
for w in c3 c2 c4close; do
  for e in RTC_X=0 RTC_POOL_LAST=1; do
    env $e timeout 300 python tools/rank_emul.py --workload $w --gpus 1 --frames 5 > gpurun_out/r2o_w.log 2>&1; echo "$w $e $(tail -n 1 gpurun_out/r2o_w.log | cut -c16-70)"
  done
done
