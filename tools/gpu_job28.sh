#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_rope.py tests/test_models.py -m gpu -x -q > gpurun_out/pytest_rope.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_rope.log
timeout 600 python tools/bench_rope.py gpurun_out/bench_rope.json 2>&1 | cut -c1-500
