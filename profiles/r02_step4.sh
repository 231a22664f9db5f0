#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step4
python -m pytest tests/test_gpu_linear.py -m gpu -x -q -k "lanes or explicit" > ${O}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a ${O}_pytest.log
tail -n 5 ${O}_pytest.log
python tools/tune_lin.py lanes:/2 lanes:/3 lanes:/4 lanes:8/3 lanes:8/4 lanes:7/4 lanes16:16 > ${O}_tune.log 2>&1; echo "tune rc=$?"
cat ${O}_tune.log
