#!/bin/bash
# the bench line of every named configuration and of the next rows (for profiles/)
mkdir -p gpurun_out
for w in c1 c2 c3 c4 c5shard c5; do
  timeout 900 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$w.log 2>&1; echo "$w rc=$?"
done
for w in f0n sampler; do
  timeout 600 python bench.py --workload $w > gpurun_out/bench_$w.log 2>&1; echo "$w rc=$?"
  timeout 600 python bench.py --workload $w --impl reference --steps 3 --warmup 1 > gpurun_out/bench_${w}_ref.log 2>&1
done
for w in c1 c2 c3 c4 c5shard c5 f0n sampler; do tail -n 1 gpurun_out/bench_$w.log | cut -c1-230; done
