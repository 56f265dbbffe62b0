#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/sampler_times.py > gpurun_out/sampler_times.log 2>&1; echo "times rc=$?"
timeout 600 python tools/sampler_times.py 64 128 5 bf16 2.0 >> gpurun_out/sampler_times.log 2>&1
timeout 600 python tools/sampler_times.py 1 128 5 bf16 >> gpurun_out/sampler_times.log 2>&1
timeout 600 python bench.py --workload sampler > gpurun_out/bench_sampler.log 2>&1; echo "bench rc=$?"
cat gpurun_out/sampler_times.log
tail -n 3 gpurun_out/bench_sampler.log
