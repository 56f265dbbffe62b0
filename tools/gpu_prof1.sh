#!/bin/bash
# one full ncu capture (with source) of a single conv_umma launch: tools/gpu_prof1.sh k,res,stats
mkdir -p gpurun_out
c=$1; n=${c//,/_}
python tools/umma_bench.py --one $c --iters 1 > gpurun_out/plain_$n.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:conv_umma -s 2 -c 1 -o gpurun_out/prof_$n -f \
    python tools/umma_bench.py --one $c --iters 1 > gpurun_out/ncu_$n.log 2>&1
echo "$c rc=$?"
