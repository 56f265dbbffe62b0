#!/bin/bash
# last check of a build: all GPU tests, smoke, the default bench line and the reference arm
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --tb=short -x > gpurun_out/t_gpu.log 2>&1; echo "tests rc=$?"; tail -n 2 gpurun_out/t_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 1 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.log 2>&1; echo "bench rc=$?"; tail -n 1 gpurun_out/bench.log | cut -c1-1200
