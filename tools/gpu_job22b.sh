#!/usr/bin/env bash
# Evidence run B: one ncu --set full capture of the forward kernel on the C3 shape.
set -u
mkdir -p gpurun_out
python tools/prof_one.py c3 --iters 3 > gpurun_out/plain22b.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fa_fwd -s 2 -c 1 -f -o gpurun_out/prof_v5_c3 python tools/prof_one.py c3 --iters 3 > gpurun_out/ncu22b.log 2>&1
echo "ncu rc=$?"
