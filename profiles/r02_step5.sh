#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step5
python -m pytest tests/test_gpu_configs.py tests/test_gpu_linear.py tests/test_gpu_parity.py tests/test_gpu_dropin.py -m gpu -x -q > ${O}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a ${O}_pytest.log
tail -n 15 ${O}_pytest.log
python tools/tune_lin.py :/@7,7 lanes16:/@7,7 warp:/@7,7 cta:/@7,7 > ${O}_tune_d7.log 2>&1; echo "tune rc=$?"
cat ${O}_tune_d7.log
python tools/e2e_probe.py > ${O}_e2e.log 2>&1; echo "e2e rc=$?"; cat ${O}_e2e.log
