#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu31.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu31.log
true
