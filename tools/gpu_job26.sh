#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_decode.py tests/test_parity.py -m gpu -x -q -k "decode" > gpurun_out/pytest_decode.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_decode.log
timeout 900 python tools/bench_decode.py --iters 100 --json gpurun_out/bench_decode.json 2>&1 | tee gpurun_out/bench_decode.log | cut -c1-900
