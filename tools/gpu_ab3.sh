#!/bin/bash
# same-box A/B of gpurun_ab/libst2dec_{old,new}.so: 3 interleaved rounds of the C2 bench (graph replay)
mkdir -p gpurun_out
for round in 1 2 3; do
for v in old new; do
  cp gpurun_ab/libst2dec_$v.so styletts2_b200/_lib/libst2dec.so
  timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ab_bench_$v$round.log 2>&1
  echo "$v$round: $(tail -n 1 gpurun_out/ab_bench_$v$round.log | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["roofline"]["kernel_ms_per_step"], d["clocks"]["sm_mhz"], [(k["us"]) for k in d["hbm_kernels"]])')"
done
done
cp gpurun_ab/libst2dec_new.so styletts2_b200/_lib/libst2dec.so
