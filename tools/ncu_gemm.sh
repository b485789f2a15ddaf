#!/bin/bash
# ncu --set full captures of three tcgen05 GEMM instantiations (forward BN=128, wgrad, dgrad); run under gpurun after a plain run
set -u
python tools/profile_step.py --steps 2 > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; exit 1; }
for spec in "0:0:128:3" "1:0:64:6" "0:1:64:6"; do
  IFS=: read a b n skip <<< "$spec"
  rx="regex:gemm_tc_kernel<\\(int\\)$a, \\(int\\)$b, \\(int\\)$n>"
  timeout 300 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "$rx" -s $skip -c 2 \
      -o gpurun_out/prof_gemm_${a}${b}${n} python tools/profile_step.py --steps 2 > gpurun_out/ncu_full_${a}${b}${n}.log 2>&1
  tail -1 gpurun_out/ncu_full_${a}${b}${n}.log
done
ls -la gpurun_out/*.ncu-rep
