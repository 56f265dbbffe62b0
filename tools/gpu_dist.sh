#!/bin/bash
# producers' prefetch distance (ST2_UMMA_DIST; default na - 2) on the current build
for d in 1 2 3; do
  echo "== dist $d"
  ST2_UMMA_DIST=$d timeout 200 python tools/umma_bench.py 2>&1 | grep -E "s2 k3 d1 c|s2 k7 d1 c|s2 k11 d1 c|s1 k7|s1 k11" | cut -c1-64
done
