#!/bin/bash
# depth of the weight ring (ST2_UMMA_WST), with and without the other roles (ST2_UMMA_DBG)
for w in 3 4 5 2; do for d in 0 30; do
  echo "== wst $w dbg $d"
  ST2_UMMA_WST=$w ST2_UMMA_DBG=$d timeout 300 python tools/umma_bench.py 2>&1 | grep -E "s2 k3 d1 c|s2 k7 d1 c|s2 k11 d1 c|s1 k7|s1 k11" | cut -c1-64
done; done
