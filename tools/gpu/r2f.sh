#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2f_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r2f_pytest.log
for g in 1 2 4 8; do
  RTC_PATHS_V1=1 timeout 300 python tools/rank_emul.py --gpus $g --frames 4 > gpurun_out/r2f_v1_n${g}.log 2>&1; echo "v1 n$g $(tail -n 1 gpurun_out/r2f_v1_n${g}.log | cut -c16-70)"
done
for v in base ns80w12 ns64w12; do
  L=raytracingc_b200/librtc_b200_$v.so; [ $v = base ] && L=raytracingc_b200/librtc_b200.so
  for g in 1 2 4 8; do
    for d in 1 2; do
      RTC_POOL_SHARE_DIV=$d RTC_B200_LIB=$L timeout 300 python tools/rank_emul.py --gpus $g --frames 4 > gpurun_out/r2f_${v}_n${g}_d${d}.log 2>&1; echo "$v n$g div$d rc=$? $(tail -n 1 gpurun_out/r2f_${v}_n${g}_d${d}.log | cut -c16-70,150-400)"
    done
  done
done
