#!/bin/bash
# one GPU-box pass: parity tests, train + inference bench lines, ncu launch list of the train step (after the plain run exited 0)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_train.json 2> gpurun_out/bench_train.err; echo "bench train rc=$?"; cat gpurun_out/bench_train.json
timeout 300 python bench.py --mode infer --steps 30 --warmup 5 > gpurun_out/bench_infer.json 2> gpurun_out/bench_infer.err; echo "bench infer rc=$?"; cat gpurun_out/bench_infer.json
timeout 120 python tools/profile_step.py --steps 3 > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_train.csv python tools/profile_step.py --steps 3 > gpurun_out/ncu_launch.log 2>&1
echo "ncu rc=$?"; grep -c '^"' gpurun_out/launches_train.csv
