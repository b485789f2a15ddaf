#!/bin/bash
N=${1:-2}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline --mode train 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$1', d['value'], d['ms_per_step'], d.get('dp_check',{}).get('replica_divergence'))"; }
N=$N
for seg in 1 2 3; do for ch in 0 2 4 8; do
  if [ $ch -eq 0 ]; then CVT_DP_SEGMENTS=$seg run "seg=$seg ch=default"; else CVT_DP_SEGMENTS=$seg NCCL_MAX_NCHANNELS=$ch run "seg=$seg ch=$ch"; fi
done; done
python bench.py --steps 30 --warmup 5 --no-cpu-baseline --mode train 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('n1', d['value'], d['ms_per_step'])"
