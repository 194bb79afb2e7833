#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --import-source on --clock-control none -k regex:render_pool -c 1 -f -o gpurun_out/r2d_pool_n1 \
   python tools/rank_emul.py --gpus 1 --frames 1 > gpurun_out/r2d_ncu1.log 2>&1; echo "ncu rc=$?"
timeout 600 ncu --set full --import-source on --clock-control none -k regex:render_pool -c 1 -f -o gpurun_out/r2d_pool_n8 \
   python tools/rank_emul.py --gpus 8 --frames 1 > gpurun_out/r2d_ncu8.log 2>&1; echo "ncu rc=$?"
