#!/bin/bash
# the no-allocation entry point: alone in a fresh process, then the whole GPU suite (it is collected last)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2n_alone.log 2>&1; echo "alone rc=$?"; tail -n 12 gpurun_out/r2n_alone.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2n_suite.log 2>&1; echo "suite rc=$?"; tail -n 5 gpurun_out/r2n_suite.log
