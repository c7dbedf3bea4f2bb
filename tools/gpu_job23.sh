#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python tools/ab.py --shapes c2,c3,c3b1 --configs 3:128 --iters 20 --rounds 7 2>&1 | tee gpurun_out/ab23.log
