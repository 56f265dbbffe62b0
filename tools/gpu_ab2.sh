#!/bin/bash
# tests on the new build, then same-box A/B of gpurun_ab/libst2dec_{old,new}.so on the C2 bench
mkdir -p gpurun_out
cp gpurun_ab/libst2dec_new.so styletts2_b200/_lib/libst2dec.so
timeout 1500 python -m pytest tests -q -m gpu --tb=short -x > gpurun_out/t_gpu.log 2>&1; echo "tests rc=$?"; tail -n 4 gpurun_out/t_gpu.log
for round in 1 2; do
for v in old new; do
  cp gpurun_ab/libst2dec_$v.so styletts2_b200/_lib/libst2dec.so
  timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/ab_bench_$v$round.log 2>&1
  echo "$v$round: $(tail -n 1 gpurun_out/ab_bench_$v$round.log | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["roofline"]["kernel_ms_per_step"], d["clocks"], [(k["us"], k["frac_of_hbm_peak"]) for k in d["hbm_kernels"]])')"
done
done
cp gpurun_ab/libst2dec_new.so styletts2_b200/_lib/libst2dec.so
python tools/kernel_times.py c2 bf16 2>&1 | grep -v -i warn | head -16
