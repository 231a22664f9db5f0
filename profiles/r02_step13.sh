#!/usr/bin/env bash
# round 2, step 13: CTA variant (d=11, cfg4) with the chain-walk read-out and one CTA-wide tally row
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step13
python -m pytest tests/test_gpu_linear.py tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_configs.py tests/test_gpu_dropin.py -m gpu -x -q > ${O}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a ${O}_pytest.log
tail -n 6 ${O}_pytest.log
python tools/tune_lin.py ':/@11,11' ':/@11,11!nb' > ${O}_tune.log 2>&1
QEC_B200_LIN_THREADS=1024 python tools/tune_lin.py ':/@11,11!nb' >> ${O}_tune.log 2>&1
QEC_B200_LIN_THREADS=256 python tools/tune_lin.py ':/@11,11!nb' >> ${O}_tune.log 2>&1
QEC_B200_LIN_STREAMS=32 python tools/tune_lin.py ':/@11,11!nb' >> ${O}_tune.log 2>&1
QEC_B200_LIN_STREAMS=128 python tools/tune_lin.py ':/@11,11!nb' >> ${O}_tune.log 2>&1
QEC_B200_LIN_CTAS=1 QEC_B200_LIN_LGB=9 python tools/tune_lin.py ':/@11,11!nb' >> ${O}_tune.log 2>&1
python tools/tune_lin.py 'cta:/@7,7' 'lanes16:/@7,7' >> ${O}_tune.log 2>&1
cat ${O}_tune.log | cut -c1-230
