#!/bin/bash
for cfg in "0 30 3" "0 31 3" "1 30 6" "1 30 8" "1 30 3"; do
  set -- $cfg
  echo "== pair $1 dbg $2 wst $3"
  ST2_UMMA_PAIR=$1 ST2_UMMA_DBG=$2 ST2_UMMA_WST=$3 timeout 300 python tools/umma_bench.py 2>&1 | grep -E "s2 k3 d1 c1|s2 k7 d1 c1|s2 k11 d1 c1" | cut -c1-64
done
