#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 900 python tools/tune.py --iters 10 --shapes c3b2,c3,c2,c3nc,c4s4k --halves 1 --emu 0 --split 128,64 2>&1 | tee gpurun_out/tune.log
