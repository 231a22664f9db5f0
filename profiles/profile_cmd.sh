#!/usr/bin/env bash
# Round-1 profiling recipe (run under gpurun): plain run first, then the launch list and one
# full capture of the dominant kernels.  Outputs land in gpurun_out/ (copied to profiles/).
set -euo pipefail
ARGS="--steps 2 --warmup 3 --shots 20000000 --no-cpu-baseline"
mkdir -p gpurun_out
python bench.py $ARGS > gpurun_out/plain.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/launches.csv python bench.py $ARGS > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_sample|k_enumerate' -s 3 -c 2 \
    -f -o gpurun_out/prof python bench.py $ARGS > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/plain.log
