#!/bin/bash
# same-box A/B of the lock-step resblock schedule (fused multi-job conv launches) against one conv per launch
mkdir -p gpurun_out; : > gpurun_out/lockstep.log
echo skip-tests >> gpurun_out/lockstep.log
for rep in 1 2; do
for v in 0 1; do
  for w in c2 c1 c4; do
    ST2_LOCKSTEP=$v timeout 600 python bench.py --workload $w --no-cpu-baseline --no-library-baseline --no-parity-paths --no-quality 2>/dev/null | tail -n 1 | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print("lockstep='$v' '$w'", d["ms_per_step"], d["value"], d["roofline"]["kernel_ms_per_step"], d["roofline"]["frac"], d["gpu_launches"], d["clocks"]["sm_mhz"], d["workspace_gb"])' >> gpurun_out/lockstep.log
  done
done
done
cat gpurun_out/lockstep.log
