#!/bin/bash
# all GPU tests + next-row bench lines (kept under gpurun_out/, copied into profiles/ by hand)
mkdir -p gpurun_out
S=gpurun_out/summary.txt; : > $S
run() { name=$1; shift; to=$1; shift; timeout $to "$@" > gpurun_out/$name.log 2>&1; echo "$name rc=$?" >> $S; }
run t_gpu 1500 python -m pytest tests -q -m gpu --tb=short -x
run bench_sampler 600 python bench.py --workload sampler
run bench_sampler_ref 600 python bench.py --workload sampler --impl reference --steps 3 --warmup 1
run bench_f0n 600 python bench.py --workload f0n
cat $S
tail -n 3 gpurun_out/t_gpu.log
tail -n 1 gpurun_out/bench_sampler.log | cut -c1-1800
tail -n 1 gpurun_out/bench_sampler_ref.log | cut -c1-400
tail -n 1 gpurun_out/bench_f0n.log | cut -c1-300
