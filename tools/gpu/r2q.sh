#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2q_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r2q_pytest.log
for g in 1 2 4 8; do
  timeout 300 python tools/rank_emul.py --gpus $g --frames 5 > gpurun_out/r2q_n${g}.log 2>&1; echo "auto n$g $(tail -n 1 gpurun_out/r2q_n${g}.log | cut -c16-70)"
done
for w in c3 c2 c4close c3s default; do
  timeout 300 python tools/rank_emul.py --workload $w --gpus 1 --frames 5 > gpurun_out/r2q_w.log 2>&1; echo "$w $(tail -n 1 gpurun_out/r2q_w.log | cut -c16-70)"
done
timeout 600 python bench.py --steps 10 --warmup 5 --no-cpu --no-modes > gpurun_out/r2q_bench_c4.json 2> gpurun_out/r2q_bench_c4.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2q_bench_c4.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','parity_fnv_ok','gpu_launches')}, 'e2e', d['e2e']['ms_per_step'], 'path', d['path_kernel']['ms'])
PY
