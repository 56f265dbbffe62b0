#!/bin/bash
# final verification of a build: all GPU tests, smoke, default bench (with cpu baseline), every configuration, next rows
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --tb=short -x > gpurun_out/t_gpu.log 2>&1; echo "tests rc=$?"; tail -n 2 gpurun_out/t_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 3 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.log 2>&1; echo "bench rc=$?"
for w in c1 c3 c4 c5shard chain; do
  timeout 900 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$w.log 2>&1; echo "$w rc=$?"
done
for w in f0n sampler; do
  timeout 600 python bench.py --workload $w > gpurun_out/bench_$w.log 2>&1; echo "$w rc=$?"
done
for w in "" _c1 _c3 _c4 _c5shard _chain _f0n _sampler; do tail -n 1 gpurun_out/bench$w.log | cut -c1-200; done
