#!/bin/bash
# one GPU: the round's measurement set -> gpurun_out/r2p/ (summaries only: the .ncu-rep files are reduced to CSV/text on the box)
O=gpurun_out/r2p; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 2 $O/pytest_gpu.log
timeout 900 python bench.py --steps 10 --warmup 5 > $O/bench_c4.json 2> $O/bench_c4.err; echo "bench c4 rc=$?"
for w in c1 c2 c3 c3s default c4close; do
  timeout 600 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu --no-modes > $O/bench_$w.json 2> $O/bench_$w.err; echo "bench $w rc=$?"
done
timeout 900 python bench.py --workload c5l --steps 2 --warmup 3 > $O/bench_c5l.json 2> $O/bench_c5l.err; echo "bench c5l rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 3 > $O/bench_reference_c4.json 2> $O/bench_reference_c4.err; echo "reference rc=$?"
# launch list of the bench command (serialised, cold cache: shares, not absolutes)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_bench_c4.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-modes > $O/launches.log 2>&1; echo "launch list rc=$?"
cap() { # name kernel-regex command...
  local name=$1 rx=$2; shift 2
  timeout 900 ncu --set full --import-source on --clock-control none -k regex:$rx -c 1 -f -o $O/$name "$@" > $O/$name.log 2>&1
  if [ -f $O/$name.ncu-rep ]; then
    ncu -i $O/$name.ncu-rep --page raw --csv > $O/$name.raw.csv 2>/dev/null
    ncu -i $O/$name.ncu-rep --page source --csv > $O/$name.src.csv 2>/dev/null
    ncu -i $O/$name.ncu-rep --page source --print-source cuda,sass --csv > $O/$name.cuda.csv 2>/dev/null
    python tools/ncu_source_summary.py $O/$name.src.csv --window 128 --top 12 > $O/$name.source_summary.txt 2>&1
    python tools/ncu_lines.py $O/$name.cuda.csv --top 40 > $O/$name.lines.txt 2>&1
    rm -f $O/$name.ncu-rep $O/$name.src.csv $O/$name.cuda.csv
    echo "captured $name"
  else echo "capture $name FAILED"; tail -n 3 $O/$name.log; fi
}
RTC_PATHS=serial cap paths_serial_c4 render_paths_kernel python tools/render_once.py --frames 1
cap pool_c4_n8share render_pool_kernel python tools/rank_emul.py --gpus 8 --frames 1
RTC_PATHS=pool cap pool_c4_full render_pool_kernel python tools/render_once.py --frames 1
cap skyblock_c4 render_skyblock_kernel python tools/render_once.py --frames 1
cap blocktask_c4 render_blocktask_kernel python tools/render_once.py --frames 1
RTC_PATHS=serial cap paths_stream_c5 render_paths_kernel python tools/c5_lattice.py --step 32 --spp 4
RTC_PATHS=pool cap pool_stream_c5 render_pool_kernel python tools/c5_lattice.py --step 32 --spp 4
# FP32 peak + issue-rate probes
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o /tmp/pipe_probe tools/pipe_probe.cu 2> $O/pipe_probe_build.log && /tmp/pipe_probe > $O/pipe_probe.txt 2>&1
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Iraytracingc_b200/csrc -Iinclude -o /tmp/sweep_probe tools/sweep_probe.cu 2> $O/sweep_probe_build.log && /tmp/sweep_probe 120 400 > $O/sweep_probe.txt 2>&1
python - > $O/fp32_peak.json <<'PY'
import json, subprocess
from raytracingc_b200 import api
vals = [api.fp32_peak_tflops(0, 5) for _ in range(3)]
clk = subprocess.run(["nvidia-smi", "--query-gpu=name,clocks.sm,clocks.max.sm,power.draw", "--format=csv,noheader"], capture_output=True, text=True).stdout.strip()
print(json.dumps({"fp32_ffma_tflops": vals, "best": max(vals), "nominal": 2 * 128 * 148 * 1.965e9 / 1e12, "how": "rtc_fp32_peak_probe: 8 independent FFMA chains per thread, 256 threads x 148*32 CTAs, best of 5, three runs", "nvidia_smi_after": clk}))
PY
# the stock 12-thread reference binary on C1 at its compiled-in 4000 spp, for the record
if [ -x oracle/_ref/dropin/rayt_stock ]; then
  ( cd oracle/_ref/dropin && echo "host cores: $(nproc); the UNMODIFIED reference program (12 pthreads, 4000 spp compiled in), C1 = simplest.obj 320x240, 1 bounce:" && TIMEFORMAT='wall %R s, user %U s, sys %S s' && time ./rayt_stock -i 3Dmodels/simplest.obj -s 320 240 -b 1 -o /tmp/stock_c1.bmp > /dev/null ) > $O/stock_rayt_c1.txt 2>&1; cat $O/stock_rayt_c1.txt
fi
python tools/c5_lattice.py --step 8 --spp 2 --flags 16 > $O/c5_lattice.txt 2>&1
RTC_TRACE=1 python tools/e2e_trace.py --frames 10 > $O/e2e_trace_c4.txt 2> $O/e2e_trace_c4.err; grep timeline $O/e2e_trace_c4.err | tail -n 3 >> $O/e2e_trace_c4.txt
ls -la $O | head -60
