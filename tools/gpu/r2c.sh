#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r2c_pytest.log
for v in base up2 w10 ns64w12; do
  L=raytracingc_b200/librtc_b200_$v.so; [ $v = base ] && L=raytracingc_b200/librtc_b200.so
  for g in 1 8; do
    RTC_B200_LIB=$L timeout 300 python tools/rank_emul.py --gpus $g --frames 4 > gpurun_out/r2c_${v}_n${g}.log 2>&1; echo "$v n$g rc=$? $(tail -n 1 gpurun_out/r2c_${v}_n${g}.log | cut -c1-400)"
  done
done
