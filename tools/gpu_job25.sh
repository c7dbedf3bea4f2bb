#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python tools/bench_decode.py --iters 100 --json gpurun_out/bench_decode.json 2>&1 | tee gpurun_out/bench_decode.log | cut -c1-600
