#!/bin/bash
# re-entry check: all GPU tests, smoke, the default bench line and the reference arm
mkdir -p gpurun_out
S=gpurun_out/summary.txt; : > $S
run() { name=$1; shift; to=$1; shift; timeout $to "$@" > gpurun_out/$name.log 2>&1; echo "$name rc=$?" >> $S; }
run t_gpu 1500 python -m pytest tests -q -m gpu --tb=short -x --durations=8
run smoke 300 python -c "import __graft_entry__ as g; g.smoke()"
run bench 900 python bench.py
run bench_c4 600 python bench.py --workload c4 --no-cpu-baseline
run bench_ref 600 python bench.py --impl reference --steps 2 --warmup 1
cat $S
tail -n 15 gpurun_out/t_gpu.log
tail -n 2 gpurun_out/smoke.log
tail -n 1 gpurun_out/bench.log | cut -c1-2500
tail -n 1 gpurun_out/bench_c4.log | cut -c1-400
tail -n 1 gpurun_out/bench_ref.log | cut -c1-600
