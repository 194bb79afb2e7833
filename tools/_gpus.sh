N=$1; O=gpurun_out/r1h; mkdir -p $O
if [ "$N" = 2 ]; then timeout 900 python -m pytest tests -m gpu -q > $O/pytest_gpu_2gpu.log 2>&1; tail -3 $O/pytest_gpu_2gpu.log; fi
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > $O/scale_n$N.json 2> $O/scale_n$N.err; echo "rc=$?"; tail -3 $O/scale_n$N.err
tail -1 $O/scale_n$N.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('N', d['n_gpus'], 'frame %.3f e2e %.3f'%(d['ms_per_step'], d['e2e']['ms_per_step']), d['e2e']['api'][:60])"
