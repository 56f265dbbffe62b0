#!/bin/bash
# same-box A/B of the resblock schedules: ST2_LOCKSTEP=0 one conv per launch, 3 three convs per launch, 1 size-dependent
# (large problems: k3 + k11 fused, k7 alone)
mkdir -p gpurun_out; : > gpurun_out/lockstep.log
for rep in 1 2; do
for v in ${LOCKSTEP_MODES:-0 3 1}; do
  for w in ${LOCKSTEP_WL:-c2 c1 c4}; do
    ST2_LOCKSTEP=$v timeout 600 python bench.py --workload $w --no-cpu-baseline --no-library-baseline --no-parity-paths --no-quality 2>/dev/null | tail -n 1 | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print("lockstep='$v' '$w'", d["ms_per_step"], d["value"], d["roofline"]["kernel_ms_per_step"], d["roofline"]["frac"], d["gpu_launches"], d["clocks"]["sm_mhz"], d["workspace_gb"])' >> gpurun_out/lockstep.log
  done
done
done
cat gpurun_out/lockstep.log
