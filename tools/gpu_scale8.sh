#!/bin/bash
# 2/4/8-GPU lines of the named shapes c3, c4, c5 on ONE 8-GPU box (tools/gpu.sh --gpus 8 -- 'bash tools/gpu_scale8.sh').
# Runs that do not need all 8 GPUs share the box side by side on disjoint GPUs (CUDA_VISIBLE_DEVICES), which cuts the
# box time by a third; every run times itself on the device (max over its ranks), so neighbours on other GPUs do not
# enter its number (they do share the host and the node's power budget: noted in the profile file).
mkdir -p gpurun_out
run() { # name, visible devices, n, port, workload, extra args
  CUDA_VISIBLE_DEVICES=$2 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $3 --master-addr 127.0.0.1 \
    --master-port $4 bench.py --gpus $3 --workload $5 --no-cpu-baseline --no-library-baseline $6 2> gpurun_out/$1.err | tail -n 1 > gpurun_out/$1.json
  echo "$1 rc=$?"
}
A="--steps 20 --warmup 3"
python -c "import torch; print(torch.cuda.device_count(), 'GPUs')"
run scale_c5_2 0,1 2 29511 c5 "" & run scale_c4_2 2,3 2 29512 c4 "$A" & run scale_c3_2 4,5 2 29513 c3 "$A" & wait
run scale_c5_4 0,1,2,3 4 29521 c5 "" & run scale_c4_4 4,5,6,7 4 29522 c4 "$A" & wait
run scale_c3_4 0,1,2,3 4 29523 c3 "$A" & run scale_c2_4 4,5,6,7 4 29524 c2 "$A" & wait
run scale_c5_8 0,1,2,3,4,5,6,7 8 29531 c5 ""
run scale_c4_8 0,1,2,3,4,5,6,7 8 29532 c4 "$A"
run scale_c3_8 0,1,2,3,4,5,6,7 8 29533 c3 "$A"
for f in gpurun_out/scale_*.json; do echo "$f: $(cut -c1-160 $f)"; done
