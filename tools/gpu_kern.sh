#!/bin/bash
# kernel unit tests + conv micro-benchmark only (fast iteration on conv_umma.cu)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -q -m gpu --tb=short -x -k "umma" > gpurun_out/t_kern.log 2>&1; echo "t_kern rc=$?"
grep -E "passed|failed|error" gpurun_out/t_kern.log | tail -2
grep -E "^E  " gpurun_out/t_kern.log | head -6
timeout 600 python tools/umma_bench.py > gpurun_out/ub.log 2>&1; cat gpurun_out/ub.log
