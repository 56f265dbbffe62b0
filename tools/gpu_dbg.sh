#!/bin/bash
# ST2_UMMA_DBG timing experiments (conv_umma.cu, UmLayout::dbg): which role bounds a tile
for d in ${DBGS:-0 2 32 4 64 128 8 0}; do
  echo "== dbg $d"
  ST2_UMMA_DBG=$d timeout 300 python tools/umma_bench.py 2>&1 | grep -E "s2 k3 d1 c1|s2 k7 d1 c1|s2 k11 d1 c1|s2 k3 lrelu" | cut -c1-64
done
