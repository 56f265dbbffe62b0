#!/bin/bash
timeout 200 python tools/pair_debug.py 2>&1 | tail -10
echo "== pair 1"
ST2_UMMA_PAIR=1 timeout 300 python tools/umma_bench.py 2>&1 | grep -E "s2 k|s1 k" | cut -c1-64
echo "== pair 1 dbg 30"
ST2_UMMA_PAIR=1 ST2_UMMA_DBG=30 timeout 300 python tools/umma_bench.py 2>&1 | grep -E "s2 k3 d1 c1|s2 k7 d1 c1|s2 k11 d1 c1" | cut -c1-64
echo "== pair 0"
ST2_UMMA_PAIR=0 timeout 300 python tools/umma_bench.py 2>&1 | grep -E "s2 k3 d1 c1|s2 k7 d1 c1|s2 k11 d1 c1" | cut -c1-64
