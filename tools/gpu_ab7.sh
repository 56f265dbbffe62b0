#!/bin/bash
# new build: kernel + decoder tests; then same-box A/B (gpurun_ab/libst2dec_{old,new}.so) of the C2 / C4 steps and the sampler
cp gpurun_ab/libst2dec_new.so styletts2_b200/_lib/libst2dec.so
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_decoder_gpu.py tests/test_sampler_gpu.py -q -m gpu --tb=short -x 2>&1 | tail -2
for v in old new old new; do
  cp gpurun_ab/libst2dec_$v.so styletts2_b200/_lib/libst2dec.so
  for w in c2 c4; do
    timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -n 1 | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print("'$v' '$w'", d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["kernel_ms_per_step"], d["roofline"]["frac"], d["clocks"]["sm_mhz"])'
  done
  timeout 600 python bench.py --workload sampler --no-cpu-baseline 2>/dev/null | tail -n 1 | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print("'$v' sampler", d["ms_per_step"])'
done
cp gpurun_ab/libst2dec_new.so styletts2_b200/_lib/libst2dec.so
