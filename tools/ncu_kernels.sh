#!/bin/bash
# ncu --set full of every launch of one kernel (regex $1) in the second train step: tools/ncu_kernels.sh <regex> <launches per step> <out name>
set -u
python tools/profile_step.py --steps 2 > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; exit 1; }
CVT_NO_GRAPH=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:$1 -s $2 -c $2 -f -o gpurun_out/prof_$3 \
    python tools/profile_step.py --steps 2 > gpurun_out/ncu_full_$3.log 2>&1
tail -2 gpurun_out/ncu_full_$3.log
