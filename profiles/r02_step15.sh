#!/usr/bin/env bash
# round 2, step 15: CTA variant defaults (1024 threads, one stream per warp): whole GPU suite,
# cfg4 timing, one full ncu capture of k_sample_lin at d=11
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step15
python -m pytest tests -m gpu -x -q > ${O}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a ${O}_pytest.log
tail -n 5 ${O}_pytest.log
python tools/tune_lin.py ':/@11,11' ':/@11,11!nb' ':/@7,7' > ${O}_tune.log 2>&1
cat ${O}_tune.log | cut -c1-230
ncu --set full --clock-control none --import-source on -k regex:'k_sample_lin' -s 2 -c 1 \
    -f -o ${O}_full_cta python tools/tune_lin.py ':/@11,11!nb' > ${O}_ncu_full.log 2>&1; echo "full rc=$?"
