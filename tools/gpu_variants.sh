#!/bin/bash
# same-box A/B of conv_umma variants: microbenchmarks + C2/C4 steps
mkdir -p gpurun_out
for v in "$@"; do
  cp gpurun_ab/libst2dec_$v.so styletts2_b200/_lib/libst2dec.so
  echo "=== $v" >> gpurun_out/variants.log
  timeout 600 python tools/umma_bench.py >> gpurun_out/variants.log 2>&1
  timeout 600 python tools/umma_bench.py --hifi 2>&1 | head -9 >> gpurun_out/variants.log
  for w in c2 c4; do
    timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline --no-library-baseline 2>/dev/null | tail -n 1 | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print("'$v' '$w'", d["ms_per_step"], d["roofline"]["kernel_ms_per_step"], d["roofline"]["frac"], d["clocks"]["sm_mhz"])' >> gpurun_out/variants.log
  done
done
cat gpurun_out/variants.log
