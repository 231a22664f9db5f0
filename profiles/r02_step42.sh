#!/usr/bin/env bash
# round 2, step 42: full capture of the current CTA variant at d=11 (rev r2e, paired record loads)
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step42
ncu --set full --clock-control none --import-source on -k regex:'k_sample_lin' -s 2 -c 1 \
    -f -o ${O}_full_cta python tools/tune_lin.py ':/@11,11!nb' > ${O}_ncu_full.log 2>&1; echo "full rc=$?"
