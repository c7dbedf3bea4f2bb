#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python tools/tune.py --iters 10 --shapes c2,c3,c3nc,c4s4k --halves 0 --emu 1,2,3 --split 128,64 2>&1 | tee gpurun_out/tune17.log
