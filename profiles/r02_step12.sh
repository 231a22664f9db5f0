#!/usr/bin/env bash
# round 2, step 12: the d=11 sampler (cfg4, CTA variant k_sample_lin): knobs and one full ncu capture
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step12
python tools/tune_lin.py ':/@11,11' ':/@11,11!nb' > ${O}_tune.log 2>&1
QEC_B200_LIN_THREADS=1024 python tools/tune_lin.py ':/@11,11!nb' >> ${O}_tune.log 2>&1
QEC_B200_LIN_THREADS=256 python tools/tune_lin.py ':/@11,11!nb' >> ${O}_tune.log 2>&1
QEC_B200_LIN_STREAMS=16 python tools/tune_lin.py ':/@11,11!nb' >> ${O}_tune.log 2>&1
QEC_B200_LIN_STREAMS=64 python tools/tune_lin.py ':/@11,11!nb' >> ${O}_tune.log 2>&1
QEC_B200_LIN_CTAS=1 QEC_B200_LIN_LGB=8 python tools/tune_lin.py ':/@11,11!nb' >> ${O}_tune.log 2>&1
QEC_B200_LIN_CTADIFF=0 python tools/tune_lin.py ':/@11,11!nb' >> ${O}_tune.log 2>&1
cat ${O}_tune.log | cut -c1-230
ncu --set full --clock-control none --import-source on -k regex:'k_sample_lin' -s 2 -c 1 \
    -f -o ${O}_full_cta python tools/tune_lin.py ':/@11,11!nb' > ${O}_ncu_full.log 2>&1; echo "full rc=$?"
