#!/bin/bash
# new build: kernel + sampler tests, sampler timing; then same-box A/B of the conv micro-benchmark and the C2 bench
mkdir -p gpurun_out
cp gpurun_ab/libst2dec_new.so styletts2_b200/_lib/libst2dec.so
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_sampler_gpu.py -q -m gpu --tb=short -x -s > gpurun_out/t_new.log 2>&1; echo "tests rc=$?"; tail -n 8 gpurun_out/t_new.log
python tools/sampler_times.py 2>&1 | grep -v -i warn | head -14
for v in old new old new; do
  cp gpurun_ab/libst2dec_$v.so styletts2_b200/_lib/libst2dec.so
  echo "== $v"; python tools/umma_bench.py 2>&1 | grep -E "s2 k3 d1 c1|s2 k7 d1 c2|s2 k11 d1 c1|s1 k7"
  timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -n 1 | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["kernel_ms_per_step"], d["clocks"]["sm_mhz"])'
done
cp gpurun_ab/libst2dec_new.so styletts2_b200/_lib/libst2dec.so
