#!/bin/bash
# Histogram of the Blackwell-specific SASS opcodes per kernel of libcvt_b200.so (tcgen05 = UTCHMMA / UTCBAR, TMEM = LDTM / STTM,
# TMA = UTMALDG / UTMASTG): the evidence that the tensor-core kernels are sm_100a-native.   tools/sass_opcodes.sh > profiles/r02_sass_opcodes.txt
LIB=${1:-transformer-stm_b200/csrc/libcvt_b200.so}
cuobjdump -sass "$LIB" | awk '
  /Function :/ { fn=$3 }
  /UTCHMMA|UTCBAR|LDTM|STTM|UTMALDG|UTMASTG|UTCCP|SYNCS/ {
    for (i=1;i<=NF;i++) if ($i ~ /^(UTCHMMA|UTCBAR|LDTM|STTM|UTMALDG|UTMASTG|UTCCP|SYNCS)/) { split($i,a,"."); c[fn" "a[1]]++ ; t[a[1]]++ }
  }
  END { for (k in c) print k, c[k] | "sort"; close("sort"); print "---- totals ----"; for (k in t) print k, t[k] }'
