#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu21.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu21.log
timeout 600 python tools/ab.py --shapes c2,c3,c3b1,c4s4k --configs 3:128,2:64 --iters 20 --rounds 5 2>&1 | tee gpurun_out/ab21.log
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline | cut -c1-200
