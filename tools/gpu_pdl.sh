#!/bin/bash
# A/B of programmatic dependent launch (ST2_PDL=0/1): tests once, then C2 and C1 benches with and without graphs
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_decoder_gpu.py tests/test_predictor_gpu.py -q -m gpu --tb=short -x 2>&1 | tail -4
for w in c2 c1; do for g in "" "--no-graphs"; do for pdl in 0 1 0 1; do
  r=$(ST2_PDL=$pdl timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline $g 2>&1 | tail -n 1 | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["e2e"]["ms_per_step"])')
  echo "$w graphs=${g:-on} pdl=$pdl: $r"
done; done; done
