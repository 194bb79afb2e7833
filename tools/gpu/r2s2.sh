#!/bin/bash
# one 8-GPU box, short form of r2s.sh (what was left of the GPU budget): C4 on 8, 4 and 2 ranks; every line checks the assembled frame
# against the golden hash (parity_fnv_ok), through the peer-write path and through render_cuda(ngpus = N)
O=gpurun_out/r2s; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
PORT=29631
run() { # N name extra-args... -> $O/$name.json
  local N=$1 name=$2; shift 2
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N "$@" > $O/$name.json 2> $O/$name.err
  echo "$name rc=$? $(tail -c 400 $O/$name.json | tr '\n' ' ' | cut -c1-200)"
  PORT=$((PORT+1))
}
run 8 scale_n8 --steps 10 --warmup 5
run 4 scale_n4 --steps 10 --warmup 5
run 2 scale_n2 --steps 10 --warmup 5
