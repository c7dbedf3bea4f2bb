#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu29.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu29.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/smoke29.log
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline | cut -c1-160
