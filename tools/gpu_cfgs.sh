#!/bin/bash
# decoder tests + the bench line of every named configuration
mkdir -p gpurun_out
S=gpurun_out/summary.txt; : > $S
run() { name=$1; shift; to=$1; shift; timeout $to "$@" > gpurun_out/$name.log 2>&1; echo "$name rc=$?" >> $S; }
run t_dec 1500 python -m pytest tests/test_decoder_gpu.py -q -m gpu --tb=short -rA -s
for w in c1 c2 c3 c4 c5shard c5; do
  run bench_$w 900 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline
done
cat $S
grep -E "passed|failed|SNR|Error|^E  " gpurun_out/t_dec.log | tail -14
for w in c1 c2 c3 c4 c5shard c5; do tail -n 1 gpurun_out/bench_$w.log | cut -c1-420; done
