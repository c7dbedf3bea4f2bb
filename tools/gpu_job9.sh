#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu9.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu9.log
for e in 0 2; do for sp in 128 96; do
timeout 120 python tools/trace_view.py c2 --halves 3 --emu $e --split $sp --dump gpurun_out/trace_c2_v4_e${e}_s$sp.json 2>&1 | tee -a gpurun_out/trace9.log
done; done
timeout 900 python tools/tune.py --iters 10 --shapes c2,c3 --halves 3 --emu 0,1,2,3 --split 128,96 2>&1 | tee gpurun_out/tune9.log
timeout 300 python tools/tune.py --iters 10 --shapes c2,c3 --halves 3 --emu 0,2 --split 64 --no-v1 2>&1 | tee -a gpurun_out/tune9.log
