#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2e_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r2e_pytest.log
for v in base ns80w12 ns64w12 ring16; do
  L=raytracingc_b200/librtc_b200_$v.so; [ $v = base ] && L=raytracingc_b200/librtc_b200.so
  for g in 1 2 8; do
    for d in 1 2; do
      RTC_POOL_SHARE_DIV=$d RTC_B200_LIB=$L timeout 300 python tools/rank_emul.py --gpus $g --frames 4 > gpurun_out/r2e_${v}_n${g}_d${d}.log 2>&1; echo "$v n$g div$d rc=$? $(tail -n 1 gpurun_out/r2e_${v}_n${g}_d${d}.log | cut -c16-70,230-400)"
    done
  done
done
