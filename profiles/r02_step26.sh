#!/usr/bin/env bash
# round 2, step 26: flush test hook, paired record loads in the CTA variant
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step26
python -m pytest tests/test_gpu_linear.py tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_configs.py tests/test_gpu_dropin.py -m gpu -x -q > ${O}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a ${O}_pytest.log
tail -n 4 ${O}_pytest.log
python tools/tune_lin.py ':/@11,11' ':/@11,11!nb' ':/@7,7' > ${O}_tune.log 2>&1
cat ${O}_tune.log | cut -c1-150
