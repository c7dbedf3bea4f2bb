#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 120 python tools/trace_view.py c2 --halves 3 --emu 0 --split 128 --dump gpurun_out/trace_c2_v4b_e0_s128.json 2>&1 | tee gpurun_out/trace10.log
timeout 120 python tools/trace_view.py c2 --halves 3 --emu 0 --split 96 --dump gpurun_out/trace_c2_v4b_e0_s96.json 2>&1 | tee -a gpurun_out/trace10.log
timeout 900 python tools/tune.py --iters 10 --shapes c2,c3 --halves 3 --emu 0,1,2 --split 128,96,64 2>&1 | tee gpurun_out/tune10.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu10.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu10.log
