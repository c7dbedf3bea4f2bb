#!/usr/bin/env bash
# Round-1 first-contact job: GPU tests, bench (both arms), ncu launch list + full capture.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.sw_power_cap,clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown --format=csv -lms 500 > gpurun_out/clocks.csv 2>/dev/null &
SMI=$!
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 30 --warmup 5 --ref-kernel > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench c3 rc=$?"
cat gpurun_out/bench_c3.json
timeout 300 python bench.py --steps 30 --warmup 5 --workload c2 --no-cpu-baseline --ref-kernel > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench c2 rc=$?"
cat gpurun_out/bench_c2.json
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
cat gpurun_out/bench_ref.json
for s in c3b1 c2 c3nc c4s4k; do timeout 120 python tools/prof_one.py $s --iters 5; done 2>&1 | tee gpurun_out/prof_one.log
kill $SMI
# ncu: launch list of the bench command, then full captures of the kernel on short commands
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fa_fwd -s 2 -c 1 -o gpurun_out/prof_c3b1 -f python tools/prof_one.py c3b1 --iters 3 > gpurun_out/ncu_c3b1.log 2>&1; echo "ncu c3b1 rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fa_fwd -s 2 -c 1 -o gpurun_out/prof_c2 -f python tools/prof_one.py c2 --iters 3 > gpurun_out/ncu_c2.log 2>&1; echo "ncu c2 rc=$?"
ls -la gpurun_out
