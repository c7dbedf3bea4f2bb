#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
python tools/prof_one.py c2 --iters 3 > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fa_fwd -s 2 -c 1 -o gpurun_out/prof_v5_c2 -f python tools/prof_one.py c2 --iters 3 > gpurun_out/ncu11.log 2>&1
echo rc=$?; cat gpurun_out/plain.log; tail -3 gpurun_out/ncu11.log
