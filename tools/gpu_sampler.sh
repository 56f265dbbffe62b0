#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sampler_gpu.py -q -m gpu --tb=short -x -s > gpurun_out/t_sampler.log 2>&1
echo "rc=$?"
tail -n 40 gpurun_out/t_sampler.log
