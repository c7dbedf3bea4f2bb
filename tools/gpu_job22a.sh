#!/usr/bin/env bash
# Evidence run A: parity tests, smoke, bench (both arms), CLI, ncu launch list.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu22.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu22.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/smoke22.log
timeout 600 python bench.py --steps 30 --warmup 5 --ref-kernel > gpurun_out/bench_v5_c3.json 2> gpurun_out/bench_v5_c3.err; echo "bench rc=$?"; cat gpurun_out/bench_v5_c3.json
timeout 600 python bench.py --steps 30 --warmup 5 --workload c2 --no-cpu-baseline --ref-kernel > gpurun_out/bench_v5_c2.json 2>> gpurun_out/bench_v5_c3.err; cat gpurun_out/bench_v5_c2.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_v5_ref.json 2>> gpurun_out/bench_v5_c3.err; cat gpurun_out/bench_v5_ref.json
timeout 600 python scripts/benchmark_kernel.py --causal --seqlen-q 8192 --seqlen-kv 8192 --batch-size 2 --dtype bfloat16 --iter 20 --no-eager --json gpurun_out/benchmark_kernel_c3b2.json 2>&1 | tail -8
timeout 600 python scripts/benchmark_kernel.py --iter 50 --json gpurun_out/benchmark_kernel_default.json 2>&1 | tail -12
timeout 600 python tools/ab.py --shapes c2,c3,c3b1,c4s4k --configs 3:128,2:64 --iters 20 --rounds 5 2>&1 | tee gpurun_out/ab22.log
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain22.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_bench_v5.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu22a.log 2>&1
echo "ncu rc=$?"
