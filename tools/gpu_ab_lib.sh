#!/bin/bash
# same-box A/B of two builds of the library (gpurun_ab/libst2dec_{old,new}.so): C2 bench + per-kernel times
mkdir -p gpurun_out
for round in 1 2; do
for v in old new; do
  cp gpurun_ab/libst2dec_$v.so styletts2_b200/_lib/libst2dec.so
  timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/ab_bench_$v$round.log 2>&1
  echo "$v$round: $(tail -n 1 gpurun_out/ab_bench_$v$round.log | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["roofline"]["kernel_ms_per_step"], d["clocks"])')"
done
done
for v in old new; do
  cp gpurun_ab/libst2dec_$v.so styletts2_b200/_lib/libst2dec.so
  timeout 600 python tools/kernel_times.py c2 bf16 --list > gpurun_out/ab_kt_$v.log 2>&1
  sed -n 17,160p gpurun_out/ab_kt_$v.log | grep conv_umma | awk '{printf "%s:%s ", $1, $3} END {print ""}'
done
