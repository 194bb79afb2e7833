#!/bin/bash
mkdir -p gpurun_out
for rep in 1 2 3; do
  for g in 8 4; do
    for e in RTC_X=0 RTC_NO_HINT=1; do
      env $e timeout 300 python tools/rank_emul.py --gpus $g --frames 8 > gpurun_out/r2r.log 2>&1; echo "rep$rep n$g $e $(tail -n 3 gpurun_out/r2r.log | cut -c16-62 | tr '\n' '|')"
    done
  done
done
