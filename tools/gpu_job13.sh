#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
for e in 0 2; do
timeout 120 python tools/trace_view.py c2 --emu $e --dump gpurun_out/trace_c2_v5_e$e.json 2>&1 | grep -v "^mma tile\|^mma P1Full" | tee -a gpurun_out/trace13.log
done
timeout 600 python tools/tune.py --iters 10 --shapes c2,c3,c4s4k --halves 0 --emu 0,1,2,3 --split 128 2>&1 | tee gpurun_out/tune13.log
