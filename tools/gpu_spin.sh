#!/bin/bash
# single-lane MMA issue loop: correctness (pair and single), micro-benchmark, kernel tests, C2 bench
timeout 200 python tools/pair_debug.py 2>&1 | tail -3
for cfg in "0 0" "1 0" "0 30" "0 14"; do
  set -- $cfg
  echo "== pair $1 dbg $2"
  ST2_UMMA_PAIR=$1 ST2_UMMA_DBG=$2 timeout 300 python tools/umma_bench.py 2>&1 | grep -E "s2 k|s1 k" | cut -c1-64
done
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_decoder_gpu.py -q -m gpu --tb=short -x 2>&1 | tail -3
for v in 0 1; do
  ST2_UMMA_PAIR=$v timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -n 1 | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print("pair '$v'", d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["kernel_ms_per_step"], d["roofline"]["frac"], d["clocks"]["sm_mhz"])'
done
