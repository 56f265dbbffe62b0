#!/bin/bash
# round profile: (1) launch list of one bench run, (2) DRAM traffic / tensor-pipe metrics of every tcgen05
# conv launch of one step (CSV, light metric set), (3) one --set full capture of a stage-2 k=11 conv
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graphs > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graphs > gpurun_out/ncu_bench.log 2>&1
python tools/profile_step.py c2 bf16 > gpurun_out/plain_step.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,sm__cycles_elapsed.max \
    --clock-control none --profile-from-start off -k regex:conv_umma -c 70 --csv --log-file gpurun_out/umma_step.csv \
    python tools/profile_step.py c2 bf16 > gpurun_out/ncu_step.log 2>&1
python tools/umma_bench.py --one 11,0,1 --iters 1 > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:conv_umma -s 2 -c 1 -o gpurun_out/prof_k11 -f \
    python tools/umma_bench.py --one 11,0,1 --iters 1 > gpurun_out/ncu_k11.log 2>&1
tail -1 gpurun_out/plain_bench.log | cut -c1-200; wc -l gpurun_out/launches.csv gpurun_out/umma_step.csv; ls -la gpurun_out
