#!/bin/bash
# 2 GPUs: multi-GPU tests, bench through torchrun (peer-write frame, render_cuda(ngpus=2) e2e), NCCL gather alternative
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2h_gpus.txt
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/r2h_pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -n 6 gpurun_out/r2h_pytest_multi.log
PORT=29511
for gm in peer nccl; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus 2 --steps 5 --warmup 3 --gather $gm > gpurun_out/r2h_bench_n2_$gm.json 2> gpurun_out/r2h_bench_n2_$gm.err; echo "bench n2 $gm rc=$?"; tail -n 2 gpurun_out/r2h_bench_n2_$gm.err
  PORT=$((PORT+1))
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2h_bench_n2_$gm.json').read().strip().splitlines()[-1])
    print({k:d[k] for k in ('value','ms_per_step','parity_fnv_ok','parity','gpu_launches')}, d['e2e'], d.get('path_kernel',{}).get('ms'))
except Exception as e: print('parse failed', e)
PY
done
RTC_TRACE=1 timeout 300 python tools/render_once.py --frames 4 2> gpurun_out/r2h_trace_n1.err > gpurun_out/r2h_trace_n1.log; grep timeline gpurun_out/r2h_trace_n1.err | tail -n 3
