#!/bin/bash
# Final evidence pass of round 2 on ONE GPU (after the 128-wide persistent attention backward kernels and the bench protocol fixes):
# tests, both bench arms, kernel-group times, ncu launch lists, ncu DRAM traffic per group, ncu --set full of the attention kernels.
# Each ncu pass runs after its plain command exited 0.
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -q > $O/r02_pytest_final.log 2>&1; echo "pytest rc=$?"; tail -2 $O/r02_pytest_final.log | cut -c1-200
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $O/r02_smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 > $O/r02_bench_final.json 2> $O/r02_bench_final.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $O/r02_bench_reference.json 2> $O/r02_bench_reference.err; echo "bench reference rc=$?"
timeout 300 python bench.py --config cfg1 --steps 20 --warmup 5 > $O/r02_bench_cfg1.json 2> $O/r02_bench_cfg1.err; echo "bench cfg1 rc=$?"
timeout 300 python bench.py --config cfg5 --steps 10 --warmup 3 --no-cpu-baseline > $O/r02_bench_cfg5_tf32.json 2> $O/r02_bench_cfg5_tf32.err; echo "bench cfg5 rc=$?"
for mode in train infer; do python tools/groups.py --mode $mode > $O/r02_groups_$mode.txt 2>&1; echo "groups $mode rc=$?"; done
timeout 120 python tools/profile_step.py --steps 3 > $O/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/r02_ncu_launches_train.csv python tools/profile_step.py --steps 3 > $O/ncu_launch.log 2>&1
echo "ncu train-step launch list rc=$?"
timeout 120 python tools/profile_step.py --steps 3 --mode infer > $O/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/r02_ncu_launches_infer.csv python tools/profile_step.py --steps 3 --mode infer > $O/ncu_launch.log 2>&1
echo "ncu infer launch list rc=$?"
for mode in train infer; do
  ncu --profile-from-start off --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
      --csv --log-file $O/traffic_$mode.csv python tools/ncu_traffic.py --run $mode --out $O/traffic_${mode}_groups.json > $O/traffic_$mode.log 2>&1
  echo "ncu traffic $mode rc=$?"
done
bash tools/ncu_infer_kernel.sh "attn_" 12 attention_train train
ls -la $O | head -40
