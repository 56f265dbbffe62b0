#!/bin/bash
# One parameterised GPU job script (replaces the one-off tools/gpu_*.sh of round 1).  Run on the GPU box through
#   tools/gpu.sh --timeout N -- 'bash tools/gpu_job.sh <job> [<job> ...]'
# Every job writes its logs under gpurun_out/ and appends "name rc=N" to gpurun_out/summary.txt.
#   tests            pytest -m gpu (all files, -x)
#   umma             tools/umma_bench.py (C2 stage shapes) + --hifi
#   waits            tools/wait_profile.py (+ --hifi): per-role barrier-wait accounting
#   bench:<wl>       bench.py --workload <wl> (no CPU baseline)
#   cap:<tag>:<case> ncu --set full --import-source capture of one conv_umma launch, <case> = umma_bench --case string
#   step:<wl>        per-launch ncu metric list (time, DRAM bytes, tensor-active) of every conv_umma launch of one step
#   launches:<wl>    ncu launch list (gpu__time_duration) of one bench run
#   sanitize         compute-sanitizer memcheck + racecheck of the kernel unit tests (small shapes)
#   scale:<wl>       bench.py --workload <wl> at N = 1 2 4 8 GPUs of one box (SCALE_NS / SCALE_ARGS override)
#   ab:<wl>          same-box A/B of gpurun_ab/libst2dec_{old,new}.so on bench workload <wl> (old new old new)
mkdir -p gpurun_out
S=gpurun_out/summary.txt
run() { name=$1; shift; to=$1; shift; timeout $to "$@" > gpurun_out/$name.log 2>&1; echo "$name rc=$?" >> $S; }
METRICS=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,sm__cycles_elapsed.max
for job in "$@"; do
  case $job in
    tests) run tests 1500 python -m pytest tests -q -m gpu -x --tb=short -s ;;
    umma) run umma 600 python tools/umma_bench.py; run umma_hifi 600 python tools/umma_bench.py --hifi ;;
    waits) run waits 600 python tools/wait_profile.py; run waits_hifi 600 python tools/wait_profile.py --hifi ;;
    bench:*) wl=${job#bench:}; run bench_$wl 900 python bench.py --workload $wl --no-cpu-baseline ;;
    cap:*) rest=${job#cap:}; tag=${rest%%:*}; cs=${rest#*:}
      run plain_$tag 300 python tools/umma_bench.py --case $cs --iters 1 &&
      run ncu_$tag 900 ncu --set full --clock-control none --import-source on -k regex:conv_umma -s 2 -c 1 \
          -o gpurun_out/prof_$tag -f python tools/umma_bench.py --case $cs --iters 1 ;;
    step:*) wl=${job#step:}
      run plain_step_$wl 600 python tools/profile_step.py $wl bf16 &&
      run ncu_step_$wl 1200 ncu --metrics $METRICS --clock-control none --profile-from-start off -k regex:conv_umma -c 200 \
          --csv --log-file gpurun_out/umma_step_$wl.csv python tools/profile_step.py $wl bf16 ;;
    launches:*) wl=${job#launches:}
      run plain_launches_$wl 600 python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline --no-graphs &&
      run ncu_launches_$wl 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
          --log-file gpurun_out/launches_$wl.csv python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline --no-graphs ;;
    sanitize)
      SEL='tests/test_kernels_gpu.py tests/test_predictor_gpu.py -k "umma or f0n_fp32 or f0n_ragged"'
      run san_mem 900 bash -c "compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest $SEL -q -m gpu -x --tb=line -p no:cacheprovider"
      run san_race 900 bash -c "compute-sanitizer --tool racecheck --error-exitcode 9 python -m pytest $SEL -q -m gpu -x --tb=line -p no:cacheprovider"
      run san_sync 900 bash -c "compute-sanitizer --tool synccheck --error-exitcode 9 python -m pytest $SEL -q -m gpu -x --tb=line -p no:cacheprovider" ;;
    scale:*) wl=${job#scale:}   # needs gpurun --gpus 8: one line per N in gpurun_out/scale_<wl>.jsonl
      for n in ${SCALE_NS:-1 2 4 8}; do
        if [ $n = 1 ]; then cmd="python bench.py"; else cmd="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py"; fi
        timeout 900 $cmd --gpus $n --workload $wl --no-cpu-baseline --no-library-baseline ${SCALE_ARGS:-} 2> gpurun_out/scale_${wl}_$n.err | tail -n 1 >> gpurun_out/scale_$wl.jsonl
        echo "scale_${wl}_$n rc=$?" >> $S
      done ;;
    ab:*) wl=${job#ab:}
      for v in old new old new; do
        cp gpurun_ab/libst2dec_$v.so styletts2_b200/_lib/libst2dec.so
        timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -n 1 | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print("'$v' '$wl'", d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["kernel_ms_per_step"], d["roofline"]["frac"], d["clocks"]["sm_mhz"])' | tee -a gpurun_out/ab_$wl.log
      done
      cp gpurun_ab/libst2dec_new.so styletts2_b200/_lib/libst2dec.so ;;
    *) echo "unknown job $job" >> $S ;;
  esac
done
cat $S
