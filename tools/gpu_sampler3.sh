#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sampler_gpu.py -q -m gpu --tb=short -x -s > gpurun_out/t_sampler.log 2>&1
echo "tests rc=$?"; tail -n 8 gpurun_out/t_sampler.log
timeout 600 python tools/sampler_times.py > gpurun_out/sampler_times.log 2>&1; echo "times rc=$?"
timeout 600 python tools/sampler_times.py 64 512 5 bf16 >> gpurun_out/sampler_times.log 2>&1
grep -v Warn gpurun_out/sampler_times.log
