#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
for cfg in "0 128" "0 64" "2 64"; do set -- $cfg; timeout 120 python tools/trace_view.py c3nc --emu $1 --split $2 --dump gpurun_out/trace_c3nc_e$1_s$2.json; done 2>&1 | tee gpurun_out/trace.log
