#!/bin/bash
# ncu --set full of the stage-1 attention kernels (forward, dK/dV, dQ) inside a train step; run under gpurun after a plain run
set -u
python tools/profile_step.py --steps 2 > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; exit 1; }
for k in attn_fwd_kernel attn_bwd_dkv_kernel attn_bwd_dq_kernel; do
  # launch order inside a step is s1, s2, s3 forward and s3, s2, s1 backward: skip to the stage-1 launch of the second step
  if [ $k = attn_fwd_kernel ]; then skip=3; else skip=5; fi
  CVT_NO_GRAPH=1 timeout 300 ncu --set full --clock-control none --import-source on -k regex:$k -s $skip -c 1 -f -o gpurun_out/prof_$k \
      python tools/profile_step.py --steps 2 > gpurun_out/ncu_full_$k.log 2>&1
  tail -1 gpurun_out/ncu_full_$k.log
done
ls -la gpurun_out/*.ncu-rep
