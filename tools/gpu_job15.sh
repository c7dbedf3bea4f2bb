#!/usr/bin/env bash
# Round-1 evidence run: parity tests, smoke, bench (both arms), ncu launch list + full capture.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu15.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu15.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/smoke15.log
timeout 600 python bench.py --steps 30 --warmup 5 --ref-kernel > gpurun_out/bench_v5_c3.json 2> gpurun_out/bench_v5_c3.err; echo "bench rc=$?"; cat gpurun_out/bench_v5_c3.json
timeout 600 python bench.py --steps 30 --warmup 5 --workload c2 --no-cpu-baseline --ref-kernel > gpurun_out/bench_v5_c2.json 2>> gpurun_out/bench_v5_c3.err; cat gpurun_out/bench_v5_c2.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_v5_ref.json 2>> gpurun_out/bench_v5_c3.err; cat gpurun_out/bench_v5_ref.json
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain15.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_bench_v5.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu15a.log 2>&1
python tools/prof_one.py c3 --iters 3 > gpurun_out/plain15b.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fa_fwd -s 2 -c 1 -f -o gpurun_out/prof_v5_c3 python tools/prof_one.py c3 --iters 3 > gpurun_out/ncu15b.log 2>&1
echo "ncu rc=$?"
