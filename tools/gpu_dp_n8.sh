#!/bin/bash
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29511 tests/scripts/dp_check.py --precision tf32 > gpurun_out/r02_dp_check_tf32_n$N.json 2> gpurun_out/r02_dp_check_tf32_n$N.err
echo "dp_check tf32 peer rc=$?"; cat gpurun_out/r02_dp_check_tf32_n$N.json; grep -i "error\|Traceback" -A3 gpurun_out/r02_dp_check_tf32_n$N.err | head -12 | cut -c1-300
show() { python -c "
import json; d=json.load(open('$1')); print('$2', round(d['value']), d['ms_per_step'], d.get('dp_check',{}).get('replica_divergence'), 'e2e', round(d['e2e']['value']), 'infer', round(d['infer']['value']) if 'infer' in d else None)"; }
$TR --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02_bench_n$N.json 2> gpurun_out/r02_bench_n$N.err
echo "bench n$N peer rc=$?"; show gpurun_out/r02_bench_n$N.json peer; grep -i "error\|Traceback" -A3 gpurun_out/r02_bench_n$N.err | head -12 | cut -c1-300
CVT_DP_NCCL=1 $TR --master-port 29514 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --mode train > gpurun_out/r02_bench_n${N}_nccl.json 2> gpurun_out/r02_bench_n${N}_nccl.err
echo "bench n$N nccl rc=$?"; show gpurun_out/r02_bench_n${N}_nccl.json nccl
CVT_DP_NCCL=1 CVT_DP_SEGMENTS=1 $TR --master-port 29515 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --mode train > gpurun_out/r02_bench_n${N}_nccl_seg1.json 2> gpurun_out/r02_bench_n${N}_nccl_seg1.err
echo "bench n$N nccl seg1 rc=$?"; show gpurun_out/r02_bench_n${N}_nccl_seg1.json nccl_seg1
