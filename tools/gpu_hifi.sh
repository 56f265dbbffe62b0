#!/bin/bash
# HiFi-GAN (config C4) check: decoder parity tests + per-kernel times
mkdir -p gpurun_out
S=gpurun_out/summary.txt; : > $S
run() { name=$1; shift; to=$1; shift; timeout $to "$@" > gpurun_out/$name.log 2>&1; echo "$name rc=$?" >> $S; }
run t_dec 1500 python -m pytest tests/test_decoder_gpu.py -q -m gpu --tb=short -rA -s
run kt4 600 python tools/kernel_times.py c4 bf16
cat $S
grep -E "passed|failed|SNR|Error|^E  " gpurun_out/t_dec.log | tail -14
head -n 14 gpurun_out/kt4.log
