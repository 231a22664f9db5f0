#!/usr/bin/env bash
# round 2, step 16: branch-free, prefetching read-out loop of the CTA variant
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step16
python -m pytest tests/test_gpu_linear.py tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_configs.py -m gpu -x -q > ${O}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a ${O}_pytest.log
tail -n 5 ${O}_pytest.log
python tools/tune_lin.py ':/@11,11' ':/@11,11!nb' ':/@7,7' > ${O}_tune.log 2>&1
cat ${O}_tune.log | cut -c1-230
