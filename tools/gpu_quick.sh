#!/bin/bash
# kernel tests, decoder tests, conv micro-bench, per-kernel times and the C2 bench
mkdir -p gpurun_out
S=gpurun_out/summary.txt; : > $S
run() { name=$1; shift; to=$1; shift; timeout $to "$@" > gpurun_out/$name.log 2>&1; echo "$name rc=$?" >> $S; }
run t_kern 900 python -m pytest tests/test_kernels_gpu.py -q -m gpu --tb=short -rA -x
if grep -q "t_kern rc=0" $S; then
run t_dec 1500 python -m pytest tests/test_decoder_gpu.py -q -m gpu --tb=short -rA -s
run ub 600 python tools/umma_bench.py
run bench 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline
run ktimes 600 python tools/kernel_times.py c2 bf16 --list
fi
cat $S
grep -E "passed|failed|error" gpurun_out/t_kern.log | tail -3
grep -E "^E  " gpurun_out/t_kern.log | head -10
grep -E "passed|failed|SNR|Error" gpurun_out/t_dec.log | tail -8
cat gpurun_out/ub.log
tail -n 2 gpurun_out/bench.log | cut -c1-300
head -n 12 gpurun_out/ktimes.log
