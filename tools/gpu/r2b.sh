#!/bin/bash
mkdir -p gpurun_out
V=raytracingc_b200/librtc_b200_v64.so
for g in 1 8; do
  RTC_B200_LIB=$V timeout 300 python tools/rank_emul.py --gpus $g --frames 4 > gpurun_out/r2b_v64_emul_n${g}.log 2>&1; echo "v64 emul $g rc=$?"; tail -n 1 gpurun_out/r2b_v64_emul_n${g}.log
  RTC_NO_SPEC=1 timeout 300 python tools/rank_emul.py --gpus $g --frames 3 > gpurun_out/r2b_nospec_emul_n${g}.log 2>&1; tail -n 1 gpurun_out/r2b_nospec_emul_n${g}.log
done
timeout 600 ncu --set full --import-source on --clock-control none -k regex:render_pool -c 1 -f -o gpurun_out/r2b_pool \
   python tools/rank_emul.py --gpus 1 --frames 1 > gpurun_out/r2b_ncu.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/
