#!/bin/bash
# first GPU session: every step has its own timeout and log; nothing aborts the rest.
mkdir -p gpurun_out
nvidia-smi > gpurun_out/smi.txt 2>&1
S=gpurun_out/summary.txt; : > $S
run() { name=$1; shift; to=$1; shift; timeout $to "$@" > gpurun_out/$name.log 2>&1; echo "$name rc=$?" >> $S; }
run t_simt 900 python -m pytest tests/test_kernels_gpu.py -q -m gpu --tb=short -rA -k "simt or source or stft or istft"
run probe 300 python tools/umma_probe.py
for i in 0 1 2 3 4 5 6 7; do
  run t_umma$i 300 python -m pytest "tests/test_kernels_gpu.py::test_conv_umma_bf16[case$i]" -q -m gpu --tb=short -rA
done
run t_umma_acc 300 python -m pytest tests/test_kernels_gpu.py -q -m gpu --tb=short -rA -k "accumulate_matches"
run t_dec 1500 python -m pytest tests/test_decoder_gpu.py -q -m gpu --tb=short -rA -s
run smoke 600 python -c "import __graft_entry__ as g; g.smoke()"
run bench_c1 600 python bench.py --steps 5 --warmup 3 --workload c1 --no-cpu-baseline
run bench 900 python bench.py --steps 5 --warmup 3
cat $S
tail -n 5 gpurun_out/bench.log
