#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu16.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu16.log
timeout 120 python tools/trace_view.py c2 --emu 2 --dump gpurun_out/trace_c2_v5b.json 2>&1 | grep -v "^mma tile\|^mma P1Full" | tee gpurun_out/trace16.log
timeout 900 python tools/tune.py --iters 10 --shapes c2,c3,c3nc,c4s4k --halves 0 --emu 1,2,3 --split 128 2>&1 | tee gpurun_out/tune16.log
