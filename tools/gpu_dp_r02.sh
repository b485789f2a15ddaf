#!/bin/bash
# N-GPU data-parallel evidence: gradient == mean of per-shard oracle gradients, replicas stay identical, bench at N ranks,
# for the peer-memory step tail (default) and for the NCCL path (CVT_DP_NCCL=1)
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for prec in fp32 tf32; do
  $TR --master-port 29511 tests/scripts/dp_check.py --precision $prec > gpurun_out/r02_dp_check_${prec}_n$N.json 2> gpurun_out/r02_dp_check_${prec}_n$N.err
  echo "dp_check $prec peer rc=$?"; cat gpurun_out/r02_dp_check_${prec}_n$N.json; grep -i "error\|Traceback" -A3 gpurun_out/r02_dp_check_${prec}_n$N.err | head -12 | cut -c1-300
done
CVT_DP_NCCL=1 $TR --master-port 29513 tests/scripts/dp_check.py --precision tf32 > gpurun_out/r02_dp_check_tf32_nccl_n$N.json 2> gpurun_out/r02_dp_check_tf32_nccl_n$N.err
echo "dp_check tf32 nccl rc=$?"; cat gpurun_out/r02_dp_check_tf32_nccl_n$N.json
show() { python -c "
import json; d=json.load(open('$1')); print('$2', round(d['value']), d['ms_per_step'], d.get('dp_check'), 'infer', round(d['infer']['value']) if 'infer' in d else None)"; }
$TR --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02_bench_n$N.json 2> gpurun_out/r02_bench_n$N.err
echo "bench n$N peer rc=$?"; show gpurun_out/r02_bench_n$N.json peer; grep -i "error\|Traceback" -A3 gpurun_out/r02_bench_n$N.err | head -12 | cut -c1-300
CVT_DP_NCCL=1 $TR --master-port 29514 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --mode train > gpurun_out/r02_bench_n${N}_nccl.json 2> gpurun_out/r02_bench_n${N}_nccl.err
echo "bench n$N nccl rc=$?"; show gpurun_out/r02_bench_n${N}_nccl.json nccl
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --mode train > gpurun_out/r02_bench_n1_ref.json 2>/dev/null; show gpurun_out/r02_bench_n1_ref.json n1
