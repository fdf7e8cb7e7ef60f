#!/bin/bash
# Blackwell-native opcodes per kernel of the built library (runs without a GPU):
#   tools/sass_opcodes.sh > profiles/rNN_sass_opcodes.txt
LIB=${1:-distributed-gpu-test_b200/b200ds/libb200ds.so}
echo "# cuobjdump -sass libb200ds.so: Blackwell-native opcodes per kernel (UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG = tensor-map TMA, UBLKCP = bulk TMA, UTCBAR = tcgen05.commit)"
cuobjdump -sass "$LIB" | awk '
  /Function :/ { fn = $3 }
  { if (match($0, /(UTC[A-Z]*MMA|LDTM|STTM|UBLKCP|UTMALDG|UTCBAR|UTCATOMSWS|SYNCS)/)) { op = substr($0, RSTART, RLENGTH); c[fn " " op]++ } }
  END { for (k in c) { split(k, a, " "); printf "%6d  %-14s %s\n", c[k], a[2], a[1] } }' | sort -k3,3 -k2,2
