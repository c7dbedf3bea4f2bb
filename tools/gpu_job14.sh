#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu14.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu14.log
rm -f gpurun_out/trace14.log
for sp in 128 96 64; do
timeout 120 python tools/trace_view.py c2 --emu 2 --split $sp --dump gpurun_out/trace_c2_v6_s$sp.json 2>&1 | grep -v "^mma tile\|^mma P1Full" | tee -a gpurun_out/trace14.log
done
timeout 900 python tools/tune.py --iters 10 --shapes c2,c3 --halves 0 --emu 0,2,3 --split 128,96,64 2>&1 | tee gpurun_out/tune14.log
timeout 300 python tools/tune.py --iters 10 --shapes c2,c3 --halves 2 --emu 2 --no-v1 2>&1 | tee -a gpurun_out/tune14.log
