#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu12.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu12.log
timeout 120 python tools/trace_view.py c2 --halves 3 --emu 8 --split 128 --dump gpurun_out/trace_c2_v5.json 2>&1 | tee gpurun_out/trace12.log
timeout 600 python tools/tune.py --iters 10 --shapes c2,c3,c3nc,c4s4k --halves 3 --emu 0,8 --split 128 2>&1 | tee gpurun_out/tune12.log
