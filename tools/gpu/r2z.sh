#!/bin/bash
# three interleaved fast deviates (RTC_NORMALS_ILP) against one looped copy
mkdir -p gpurun_out
for d in "-DRTC_NORMALS_LOOP=1" ""; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 $d -Iraytracingc_b200/csrc -Iinclude -o /tmp/shade_probe tools/shade_probe.cu 2> gpurun_out/r2z_shade_probe_build.log && timeout 120 /tmp/shade_probe | grep "random_direction\|fast path"
done
for v in base ilp base ilp; do
  L=raytracingc_b200/librtc_b200_$v.so; [ $v = base ] && L=raytracingc_b200/librtc_b200.so
  for g in 1 8; do
    RTC_B200_LIB=$L timeout 300 python tools/rank_emul.py --gpus $g --frames 6 > gpurun_out/r2z_${v}_n$g.log 2>&1; echo "$v n$g $(tail -n 2 gpurun_out/r2z_${v}_n$g.log | cut -c16-62 | tr '\n' '|')"
  done
done
RTC_B200_LIB=raytracingc_b200/librtc_b200_ilp.so timeout 600 python -m pytest tests/test_gpu_math.py tests/test_gpu_parity.py -x -q 2>&1 | tail -n 2
