#!/bin/bash
# Round-2 evidence run (one GPU): tests, sanitizer passes, ncu traffic per kernel group, the bench lines of cfg 1 / cfg 2 (batch 64 and 128) / cfg 4.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r02_pytest_c.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02_pytest_c.log | cut -c1-200
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_cfg2_b64.json 2> gpurun_out/r02_bench_cfg2_b64.err; echo "bench cfg2 rc=$?"
python bench.py --steps 20 --warmup 5 --batch 128 --no-cpu-baseline > gpurun_out/r02_bench_cfg2_b128.json 2> gpurun_out/r02_bench_cfg2_b128.err; echo "bench b128 rc=$?"
python bench.py --steps 20 --warmup 5 --config cfg1 > gpurun_out/r02_bench_cfg1.json 2> gpurun_out/r02_bench_cfg1.err; echo "bench cfg1 rc=$?"
python bench.py --steps 10 --warmup 3 --sweep --no-cpu-baseline > gpurun_out/r02_bench_cfg4_sweep_n1.json 2> gpurun_out/r02_bench_cfg4_sweep_n1.err; echo "sweep rc=$?"
for mode in train infer; do
  ncu --profile-from-start off --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
      --csv --log-file gpurun_out/traffic_$mode.csv python tools/ncu_traffic.py --run $mode --out gpurun_out/traffic_${mode}_groups.json > gpurun_out/traffic_$mode.log 2>&1
  echo "ncu traffic $mode rc=$?"
done
for tool in racecheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python -m pytest tests/test_gpu_attention.py tests/test_gpu_gemm.py -q -x -k "not many_ctas" > gpurun_out/r02_sanitizer_$tool.log 2>&1
  echo "sanitizer $tool rc=$?"; tail -3 gpurun_out/r02_sanitizer_$tool.log | cut -c1-200
done
