#!/usr/bin/env bash
# round 2, step 22: full ncu capture of the dominant kernel (k_sample_lanes, rev r2d) + launch list
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step22
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > ${O}_bench.log 2> ${O}_bench.err; echo "bench rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file ${O}_launches.csv \
    python bench.py --steps 5 --warmup 3 --no-cpu-baseline > ${O}_ncu_launches.log 2>&1; echo "launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_sample_lanes' -s 3 -c 1 \
    -f -o ${O}_full_lanes python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-enum --no-frame --plugin-shots 0 > ${O}_ncu_full.log 2>&1; echo "full rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_enum_lanes' -s 3 -c 1 \
    -f -o ${O}_full_enum python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-frame --plugin-shots 0 > ${O}_ncu_full_enum.log 2>&1; echo "full enum rc=$?"
