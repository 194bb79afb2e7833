#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2g_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/r2g_pytest.log
for g in 1 2 4 8; do
  timeout 300 python tools/rank_emul.py --gpus $g --frames 4 > gpurun_out/r2g_auto_n${g}.log 2>&1; echo "auto n$g $(tail -n 1 gpurun_out/r2g_auto_n${g}.log | cut -c16-70,150-300)"
done
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r2g_bench_c4.json 2> gpurun_out/r2g_bench_c4.err; echo "bench c4 rc=$?"; tail -n 3 gpurun_out/r2g_bench_c4.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/r2g_bench_c4.json').read().strip().splitlines()[-1])
    print({k:d[k] for k in ('value','ms_per_step','parity_fnv_ok','gpu_launches')}, d['e2e'], d.get('path_kernel'), d['roofline']['frac'], d.get('modes'))
except Exception as e: print('parse failed', e)
PY
timeout 900 python bench.py --steps 3 --warmup 3 --workload c5l --no-cpu > gpurun_out/r2g_bench_c5l.json 2> gpurun_out/r2g_bench_c5l.err; echo "bench c5l rc=$?"; tail -n 3 gpurun_out/r2g_bench_c5l.err; cut -c1-300 gpurun_out/r2g_bench_c5l.json
