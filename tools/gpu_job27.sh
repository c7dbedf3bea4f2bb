#!/usr/bin/env bash
# ncu --set full capture of the decode kernel (one shape), after the same command ran clean.
set -u
mkdir -p gpurun_out
python tools/prof_decode.py b16_s8k --iters 3 > gpurun_out/plain27.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fa_decode -s 2 -c 1 -f -o gpurun_out/prof_decode_b16_s8k python tools/prof_decode.py b16_s8k --iters 3 > gpurun_out/ncu27.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/plain27.log | tail -1
