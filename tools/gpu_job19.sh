#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu19.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu19.log
timeout 600 python scripts/benchmark_kernel.py --json gpurun_out/benchmark_kernel_default.json 2>&1 | tail -14
timeout 600 python scripts/benchmark_kernel.py --causal --dtype bfloat16 --seqlen-q 8192 --seqlen-kv 8192 --batch-size 2 --num-heads-q 32 --iter 20 --json gpurun_out/benchmark_kernel_c3b2.json 2>&1 | grep "Custom\|Official\|max-abs"
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline | cut -c1-400
