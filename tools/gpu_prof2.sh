#!/bin/bash
# full ncu captures (with source) of single conv_umma launches: k3 c1 and k11 c1 at the stage-2 shape
mkdir -p gpurun_out
for c in 3,0,1 11,0,1; do
  n=${c//,/_}
  python tools/umma_bench.py --one $c --iters 1 > gpurun_out/plain_$n.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:conv_umma -s 2 -c 1 -o gpurun_out/prof_$n -f \
      python tools/umma_bench.py --one $c --iters 1 > gpurun_out/ncu_$n.log 2>&1
  echo "$c rc=$?"
done
ls -la gpurun_out | grep prof_
