#!/bin/bash
# first GPU contact of the pool kernel: smoke, parity tests, kernel times against the first-form kernel
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv > gpurun_out/r2a_gpu.txt 2>&1
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_smoke.log 2>&1; echo "smoke rc=$?"
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2a_pytest.log
for g in 1 8; do
  timeout 300 python tools/rank_emul.py --gpus $g --frames 4 > gpurun_out/r2a_emul_n${g}.log 2>&1; echo "emul $g rc=$?"
  RTC_PATHS_V1=1 timeout 300 python tools/rank_emul.py --gpus $g --frames 4 > gpurun_out/r2a_emul_n${g}_v1.log 2>&1
done
tail -2 gpurun_out/r2a_emul_n*.log
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?"
cut -c1-600 gpurun_out/r2a_bench.json
