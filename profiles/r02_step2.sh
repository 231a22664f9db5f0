#!/usr/bin/env bash
# round 2, step 2: the half-word lane-stream sampler (k_sample_lanes16): parity tests, then
# kernel-only timing against the 32-shot lanes kernel at several warp counts.
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step2
python -m pytest tests/test_gpu_linear.py tests/test_gpu_parity.py tests/test_gpu_configs.py -m gpu -x -q > ${O}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a ${O}_pytest.log
tail -n 30 ${O}_pytest.log
python tools/tune_lin.py > ${O}_tune.log 2>&1; echo "tune rc=$?"
cat ${O}_tune.log
