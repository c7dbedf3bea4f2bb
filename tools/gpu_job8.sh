#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv | tee gpurun_out/smi.log
for sp in 128 96 64; do
timeout 120 python tools/trace_view.py c2 --halves 1 --emu 0 --split $sp --dump gpurun_out/trace_c2_h1_s$sp.json 2>&1 | tee -a gpurun_out/trace8.log
done
timeout 120 python tools/trace_view.py c3nc --halves 1 --emu 0 --split 128 2>&1 | tee -a gpurun_out/trace8.log
timeout 120 python tools/trace_view.py c3nc --halves 1 --emu 1 --split 128 2>&1 | tee -a gpurun_out/trace8.log
timeout 600 python tools/tune.py --iters 10 --shapes c2,c3 --halves 1 --emu 0,1 --split 128,96,64 2>&1 | tee gpurun_out/tune8.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu8.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu8.log
