#!/bin/bash
# kernel tests, decoder tests, then the C2 bench (each with its own timeout and log)
mkdir -p gpurun_out
S=gpurun_out/summary.txt; : > $S
run() { name=$1; shift; to=$1; shift; timeout $to "$@" > gpurun_out/$name.log 2>&1; echo "$name rc=$?" >> $S; }
run t_kern 900 python -m pytest tests/test_kernels_gpu.py -q -m gpu --tb=short -rA -x
run t_dec 1500 python -m pytest tests/test_decoder_gpu.py -q -m gpu --tb=short -rA -s
run bench 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline
run ktimes 600 python tools/kernel_times.py c2 bf16 --list
cat $S
grep -E "passed|failed|error" gpurun_out/t_kern.log | tail -3
grep -E "passed|failed|SNR|Error" gpurun_out/t_dec.log | tail -8
tail -n 2 gpurun_out/bench.log
head -n 24 gpurun_out/ktimes.log
