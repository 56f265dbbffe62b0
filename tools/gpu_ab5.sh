#!/bin/bash
# same-box A/B of two builds (gpurun_ab/libst2dec_{old,new}.so): conv tests on the new one, conv micro-benchmark and C2 bench on both
mkdir -p gpurun_out
cp gpurun_ab/libst2dec_new.so styletts2_b200/_lib/libst2dec.so
timeout 900 python -m pytest tests/test_kernels_gpu.py -q -m gpu --tb=short -x > gpurun_out/t_new.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/t_new.log
for v in old new old new; do
  cp gpurun_ab/libst2dec_$v.so styletts2_b200/_lib/libst2dec.so
  echo "== $v"; python tools/umma_bench.py 2>&1 | grep -E "s2 k3 |s2 k7 |s2 k11 |s1 k7|s1 k3"
  timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -n 1 | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["kernel_ms_per_step"], d["clocks"]["sm_mhz"])'
done
cp gpurun_ab/libst2dec_new.so styletts2_b200/_lib/libst2dec.so
