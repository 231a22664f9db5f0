#!/usr/bin/env bash
# Round-1 profiling recipe (run under gpurun): plain run first, then the launch list of the
# same command and one full capture of the dominant kernel.  Outputs land in gpurun_out/
# (copied to profiles/r01/ afterwards).
set -u
ARGS="--steps 5 --warmup 3"
mkdir -p gpurun_out
python bench.py $ARGS > gpurun_out/bench_final.log 2> gpurun_out/bench_final.err
python bench.py $ARGS --impl reference > gpurun_out/bench_reference.log 2>> gpurun_out/bench_final.err
ncu --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/launches_final.csv python bench.py $ARGS --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_sample_l' -s 3 -c 1 \
    -f -o gpurun_out/full_sampler python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-enum > gpurun_out/ncu_full.log 2>&1
tail -c 2200 gpurun_out/bench_final.log; echo; tail -c 700 gpurun_out/bench_reference.log
