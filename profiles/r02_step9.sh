#!/usr/bin/env bash
# round 2, step 9: where the lane-stream sampler stands (lanes / lanes16 / candidates per trip /
# tallies off / bits off), then the launch list of the bench and one full ncu capture of the
# dominant kernel (both after the plain run has exited 0).
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step9
python tools/tune_lin.py lanes: lanes:/1 lanes:/3 lanes:/4 'lanes:!nt' 'lanes:!nb' lanes:8 lanes:7 \
   lanes16:16 lanes16:14 lanes16:12 'lanes16:16!nt' 'lanes16:16!nb' > ${O}_tune.log 2>&1; echo "tune rc=$?"
cat ${O}_tune.log
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > ${O}_bench.log 2> ${O}_bench.err; echo "bench rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file ${O}_launches.csv \
    python bench.py --steps 5 --warmup 3 --no-cpu-baseline > ${O}_ncu_launches.log 2>&1; echo "launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_sample_lanes' -s 3 -c 1 \
    -f -o ${O}_full_lanes python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-enum --no-frame --plugin-shots 0 > ${O}_ncu_full.log 2>&1; echo "full rc=$?"
ls -la gpurun_out | tail -n 12
