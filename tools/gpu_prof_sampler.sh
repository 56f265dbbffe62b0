#!/bin/bash
# ncu evidence for the style sampler: launch list of one sampling run + full captures of the attention and AdaLN kernels
mkdir -p gpurun_out
python tools/sampler_times.py > gpurun_out/plain_sampler.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active \
    --clock-control none -c 400 --csv --log-file gpurun_out/sampler_launches.csv python tools/sampler_times.py > gpurun_out/ncu_sampler.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:attention_mma -s 30 -c 1 -o gpurun_out/prof_attn -f \
    python tools/sampler_times.py > gpurun_out/ncu_attn.log 2>&1
ncu -i gpurun_out/prof_attn.ncu-rep --page details > gpurun_out/prof_attn_details.txt 2>&1
ncu --set full --clock-control none --import-source on -k regex:adaln2 -s 30 -c 1 -o gpurun_out/prof_adaln -f \
    python tools/sampler_times.py > gpurun_out/ncu_adaln.log 2>&1
ncu -i gpurun_out/prof_adaln.ncu-rep --page details > gpurun_out/prof_adaln_details.txt 2>&1
wc -l gpurun_out/sampler_launches.csv; ls -la gpurun_out | tail -12
