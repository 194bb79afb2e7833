#!/bin/bash
# second version of the fast normal deviate: instruction costs, exhaustive self-check, A/B against the exact-only build
mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Iraytracingc_b200/csrc -Iinclude -o /tmp/shade_probe tools/shade_probe.cu 2> gpurun_out/r2x_shade_probe_build.log && timeout 120 /tmp/shade_probe > gpurun_out/r2x_shade_probe.txt 2>&1; echo "probe rc=$?"; cat gpurun_out/r2x_shade_probe.txt
timeout 600 python tools/normal_selfcheck.py > gpurun_out/r2x_normal_selfcheck.txt 2>&1; echo "selfcheck rc=$?"; tail -n 1 gpurun_out/r2x_normal_selfcheck.txt
for v in base exact base exact; do
  L=raytracingc_b200/librtc_b200_$v.so; [ $v = base ] && L=raytracingc_b200/librtc_b200.so
  for g in 1 8; do
    RTC_B200_LIB=$L timeout 300 python tools/rank_emul.py --gpus $g --frames 6 > gpurun_out/r2x_${v}_n$g.log 2>&1; echo "$v n$g $(tail -n 2 gpurun_out/r2x_${v}_n$g.log | cut -c16-62 | tr '\n' '|')"
  done
done
