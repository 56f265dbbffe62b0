#!/bin/bash
mkdir -p gpurun_out
for d in 0 2 0 2; do echo "== ST2_UMMA_DIST=$d"; ST2_UMMA_DIST=$d python tools/umma_bench.py 2>&1 | grep -E "k3 d1 c1|k3 d1 c2|k11 d1 c1|k11 d1 c2|k7 d1 c2|s1 k7"; done
