#!/bin/bash
mkdir -p gpurun_out
for v in base whatif; do
  L=raytracingc_b200/librtc_b200_$v.so; [ $v = base ] && L=raytracingc_b200/librtc_b200.so
  for g in 1 8; do
    RTC_B200_LIB=$L timeout 300 python tools/rank_emul.py --gpus $g --frames 5 > gpurun_out/r2u_${v}_n$g.log 2>&1; echo "$v n$g $(tail -n 1 gpurun_out/r2u_${v}_n$g.log | cut -c16-100)"
  done
done
