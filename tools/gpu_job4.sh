#!/usr/bin/env bash
# two-threads-per-row softmax bring-up: quick check, tests, sweep, trace
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 900 python tools/tune.py --iters 10 2>&1 | tee gpurun_out/tune.log
for cfg in "2 0" "2 1" "1 0"; do set -- $cfg; timeout 120 python tools/trace_view.py c3nc --halves $1 --emu $2 --dump gpurun_out/trace_c3nc_h$1_e$2.json; done 2>&1 | tee gpurun_out/trace.log
