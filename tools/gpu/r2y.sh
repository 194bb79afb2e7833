#!/bin/bash
# glibc's powf restated + second fast normal deviate: instruction costs, exhaustive self-check, the math tests, A/B
mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Iraytracingc_b200/csrc -Iinclude -o /tmp/shade_probe tools/shade_probe.cu 2> gpurun_out/r2y_shade_probe_build.log && timeout 120 /tmp/shade_probe > gpurun_out/r2y_shade_probe.txt 2>&1; echo "probe rc=$?"; head -12 gpurun_out/r2y_shade_probe.txt
timeout 600 python tools/normal_selfcheck.py > gpurun_out/r2y_normal_selfcheck.txt 2>&1; echo "selfcheck rc=$?"; tail -n 1 gpurun_out/r2y_normal_selfcheck.txt
timeout 600 python -m pytest tests/test_gpu_math.py tests/test_gpu_parity.py -x -q > gpurun_out/r2y_pytest_math.log 2>&1; echo "pytest math rc=$?"; tail -n 3 gpurun_out/r2y_pytest_math.log
for v in base exact pow1 base pow1; do
  L=raytracingc_b200/librtc_b200_$v.so; [ $v = base ] && L=raytracingc_b200/librtc_b200.so
  for g in 1 8; do
    RTC_B200_LIB=$L timeout 300 python tools/rank_emul.py --gpus $g --frames 6 > gpurun_out/r2y_${v}_n$g.log 2>&1; echo "$v n$g $(tail -n 2 gpurun_out/r2y_${v}_n$g.log | cut -c16-62 | tr '\n' '|')"
  done
done
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2y_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r2y_pytest.log
