#!/bin/bash
# one 8-GPU box: C4 on 2/4/8 ranks (torchrun; peer-write frame, render_cuda(ngpus=N) e2e), the NCCL-gather alternative at 8,
# the sample split of the synthetic 1 M-triangle lattice at 8, and the multi-GPU tests for 2/4/8 devices
O=gpurun_out/r2s; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q > $O/pytest_multi_8gpu.log 2>&1; echo "pytest multi rc=$?"; tail -n 3 $O/pytest_multi_8gpu.log
PORT=29531
run() { # N extra-args... -> $O/$name.json
  local N=$1 name=$2; shift 2
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N "$@" > $O/$name.json 2> $O/$name.err
  echo "$name rc=$? $(python - <<PY
import json
try:
    d=json.loads(open('$O/$name.json').read().strip().splitlines()[-1])
    print({k:d[k] for k in ('value','ms_per_step','parity_fnv_ok')}, 'e2e ms', d['e2e']['ms_per_step'], 'path ms', d.get('path_kernel',{}).get('ms'), 'prim ms', d.get('primary_kernels',{}).get('ms'))
except Exception as e: print('parse failed', e)
PY
)"
  PORT=$((PORT+1))
}
timeout 600 python bench.py --steps 10 --warmup 5 --no-cpu --no-modes > $O/scale_n1.json 2> $O/scale_n1.err; echo "n1 rc=$?"
for N in 2 4 8; do run $N scale_n$N --steps 10 --warmup 5; done
run 8 scale_n8_nccl --steps 10 --warmup 5 --gather nccl
run 8 c5l_n8_samples --workload c5l --steps 2 --warmup 3
tail -n 3 $O/*.err | tail -n 30
