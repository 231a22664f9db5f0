#!/usr/bin/env bash
# round 2, step 11+: read-out tallies in registers (unrolled trips), planes carries per trip
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step11
python -m pytest tests/test_gpu_linear.py tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_configs.py -m gpu -x -q > ${O}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a ${O}_pytest.log
tail -n 6 ${O}_pytest.log
python tools/tune_lin.py lanes: 'lanes:!nt' 'lanes:!nb' lanes:8 > ${O}_tune.log 2>&1; echo "tune rc=$?"
cat ${O}_tune.log
