#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/r2i_pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -n 3 gpurun_out/r2i_pytest_multi.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2i_bench_n2.json 2> gpurun_out/r2i_bench_n2.err; echo "bench n2 rc=$?"; tail -n 2 gpurun_out/r2i_bench_n2.err
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2i_bench_n2.json').read().strip().splitlines()[-1])
    print({k:d[k] for k in ('value','ms_per_step','parity_fnv_ok','gpu_launches')}, d['e2e'], d.get('path_kernel',{}).get('ms'))
except Exception as e: print('parse failed', e)
PY
RTC_NO_THREADS=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2i_bench_n2_nothreads.json 2> gpurun_out/r2i_bench_n2_nothreads.err; echo "bench n2 nothreads rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2i_bench_n2_nothreads.json').read().strip().splitlines()[-1])
    print('nothreads', d['e2e']['ms_per_step'])
except Exception as e: print('parse failed', e)
PY
