#!/bin/bash
# Round-2 final evidence pass on ONE GPU: tests, the driver's bench lines (both arms), cfg 1 / cfg 4 sweep, kernel-group times, ncu launch list
# of the same bench command, ncu DRAM traffic per group, ncu --set full of the kernels VERDICT names.  Each ncu pass runs after its plain command exited 0.
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -q > $O/r02_pytest_final.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r02_pytest_final.log | cut -c1-200
timeout 600 python bench.py --steps 20 --warmup 5 > $O/r02_bench_final.json 2> $O/r02_bench_final.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $O/r02_bench_reference.json 2> $O/r02_bench_reference.err; echo "bench reference rc=$?"
timeout 300 python bench.py --config cfg1 --steps 20 --warmup 5 > $O/r02_bench_cfg1.json 2> $O/r02_bench_cfg1.err; echo "bench cfg1 rc=$?"
timeout 600 python bench.py --sweep --steps 10 --warmup 3 --no-cpu-baseline > $O/r02_bench_cfg4_sweep_n1.json 2> $O/r02_bench_cfg4_sweep_n1.err; echo "sweep rc=$?"
for mode in train infer; do python tools/groups.py --mode $mode > $O/r02_groups_$mode.txt 2>&1; echo "groups $mode rc=$?"; done
# launch list of the bench command itself (graphs off so that ncu sees kernels, not graph launches)
CVT_NO_GRAPH=1 timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-parity --no-batch128 > $O/plain_bench.json 2>/dev/null && \
CVT_NO_GRAPH=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/r02_ncu_launches_bench.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-parity --no-batch128 > $O/ncu_bench.log 2>&1
echo "ncu launch list rc=$?"; grep -c '^"' $O/r02_ncu_launches_bench.csv
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
bash tools/ncu_infer_kernel.sh "attn_" 9 attention_train train
bash tools/ncu_infer_kernel.sh "mlp_fused" 2 mlp_fused_train train
bash tools/ncu_infer_kernel.sh "mlp_fused" 2 mlp_fused_infer infer
bash tools/ncu_infer_kernel.sh "gemm_tma" 12 gemm_tma_infer infer
ls -la $O | head -60
