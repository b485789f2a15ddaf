#!/bin/bash
# ncu --set full of the launches of one kernel (regex $1) in the second eager inference forward; summary as JSON lines + source hot spots
# tools/ncu_infer_kernel.sh <regex> <count> <out name> [mode]
set -u
MODE=${4:-infer}
CVT_NO_GRAPH=1 python tools/profile_step.py --steps 2 --mode $MODE > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
CVT_NO_GRAPH=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:$1 -s $2 -c $2 -f -o gpurun_out/prof_$3 \
    python tools/profile_step.py --steps 2 --mode $MODE > gpurun_out/ncu_full_$3.log 2>&1
tail -2 gpurun_out/ncu_full_$3.log
python tools/ncu_export.py gpurun_out/prof_$3.ncu-rep > gpurun_out/ncu_full_$3.jsonl
ncu -i gpurun_out/prof_$3.ncu-rep --page details --csv 2>/dev/null | grep -i "stall\|Warp Cycles Per\|No Eligible\|Issued Warp\|Theoretical Occ\|Achieved Occ\|Registers\|Shared Memory Config\|Duration" | cut -c1-260 > gpurun_out/ncu_details_$3.txt
rm -f gpurun_out/prof_$3.ncu-rep
