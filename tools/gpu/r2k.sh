#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pool.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2k_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r2k_pytest.log
for g in 1 2 4 8; do
  RTC_PATHS=serial timeout 300 python tools/rank_emul.py --gpus $g --frames 4 > gpurun_out/r2k_serial_n${g}.log 2>&1; echo "serial n$g $(tail -n 1 gpurun_out/r2k_serial_n${g}.log | cut -c16-70)"
  for v in base batch1 batch32; do
    L=raytracingc_b200/librtc_b200_$v.so; [ $v = base ] && L=raytracingc_b200/librtc_b200.so
    RTC_PATHS=pool RTC_B200_LIB=$L timeout 300 python tools/rank_emul.py --gpus $g --frames 4 > gpurun_out/r2k_${v}_n${g}.log 2>&1; echo "pool $v n$g rc=$? $(tail -n 1 gpurun_out/r2k_${v}_n${g}.log | cut -c16-70,150-400)"
  done
done
