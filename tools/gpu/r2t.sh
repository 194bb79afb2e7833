#!/bin/bash
mkdir -p gpurun_out
run() { local name=$1; shift
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2t_$name.json 2> gpurun_out/r2t_$name.err; echo "$name rc=$?"
  grep "falling back" gpurun_out/r2t_$name.err | head -n 2
  python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/r2t_$name.json').read().strip().splitlines() if l.startswith('{')][-1])
    print({k:d[k] for k in ('value','ms_per_step','parity_fnv_ok','gpu_launches')}, d['config']['split'][:60], 'e2e', round(d['e2e']['ms_per_step'],3))
except Exception as e: print('parse failed', e)
PY
  PORT=$((PORT+1)); }
PORT=29561
run peer RTC_X=0
run fallback RTC_BENCH_FORCE_IPC_FAIL=1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29570 bench.py --impl reference --gpus 2 --steps 1 --warmup 3 > gpurun_out/r2t_ref.json 2> gpurun_out/r2t_ref.err; echo "reference arm under torchrun rc=$?"; cut -c1-200 gpurun_out/r2t_ref.json
