#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/normal_selfcheck.py > gpurun_out/r2v_normal_selfcheck.txt 2>&1; echo "selfcheck rc=$?"; tail -n 2 gpurun_out/r2v_normal_selfcheck.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2v_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r2v_pytest.log
for g in 1 2 4 8; do
  timeout 300 python tools/rank_emul.py --gpus $g --frames 5 > gpurun_out/r2v_n$g.log 2>&1; echo "n$g $(tail -n 1 gpurun_out/r2v_n$g.log | cut -c16-100)"
done
for w in c3 c2 c4close c3s default; do
  timeout 300 python tools/rank_emul.py --workload $w --gpus 1 --frames 5 > gpurun_out/r2v_w.log 2>&1; echo "$w $(tail -n 1 gpurun_out/r2v_w.log | cut -c16-70)"
done
