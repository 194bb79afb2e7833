#!/bin/bash
mkdir -p gpurun_out
for c in 16 32 64 8; do
  RTC_COPY_CHUNKS=$c RTC_TRACE=1 timeout 300 python tools/e2e_trace.py --frames 10 > gpurun_out/r2m_chunks$c.log 2> gpurun_out/r2m_chunks$c.err; echo "chunks $c: $(tail -n 1 gpurun_out/r2m_chunks$c.log)"; grep timeline gpurun_out/r2m_chunks$c.err | tail -n 2
done
RTC_NO_PIPELINE=1 RTC_TRACE=1 timeout 300 python tools/e2e_trace.py --frames 10 > gpurun_out/r2m_nopipe.log 2> gpurun_out/r2m_nopipe.err; echo "no pipeline: $(tail -n 1 gpurun_out/r2m_nopipe.log)"
tail -n 4 gpurun_out/r2m_chunks16.log
timeout 600 python -m pytest tests/test_gpu_host.py tests/test_gpu_quick.py tests/test_gpu_full.py -m gpu -x -q 2>&1 | tail -n 2
