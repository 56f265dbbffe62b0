#!/bin/bash
mkdir -p gpurun_out
S=gpurun_out/summary.txt; : > $S
run() { name=$1; shift; to=$1; shift; timeout $to "$@" > gpurun_out/$name.log 2>&1; echo "$name rc=$?" >> $S; }
run t_kern 900 python -m pytest tests/test_kernels_gpu.py -q -m gpu --tb=short -rA
run t_dec 1500 python -m pytest tests/test_decoder_gpu.py -q -m gpu --tb=short -rA -s
run bench_c1 600 python bench.py --steps 5 --warmup 3 --workload c1 --no-cpu-baseline
run bench 900 python bench.py --steps 5 --warmup 3
run bench_ref 600 python bench.py --impl reference --steps 2 --warmup 1
cat $S
tail -n 3 gpurun_out/bench_c1.log; tail -n 3 gpurun_out/bench.log; tail -n 2 gpurun_out/bench_ref.log
