#!/bin/bash
# depth of the weight ring (ST2_UMMA_WST) on the current build
for w in 3 4 5; do
  echo "== wst $w"
  ST2_UMMA_WST=$w timeout 300 python tools/umma_bench.py 2>&1 | grep -E "s2 k3 d1 c|s2 k7 d1 c|s2 k11 d|s1 k" | cut -c1-64
done
