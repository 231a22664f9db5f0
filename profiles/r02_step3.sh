#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step3
python -m pytest tests/test_gpu_linear.py tests/test_gpu_parity.py -m gpu -x -q > ${O}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a ${O}_pytest.log
tail -n 12 ${O}_pytest.log
python tools/tune_lin.py lanes: lanes16:16 lanes16:12 > ${O}_tune.log 2>&1; echo "tune rc=$?"
cat ${O}_tune.log
QEC_B200_LIN_MODE=lanes16 ncu --set full --clock-control none --import-source on -k regex:'k_sample_lanes16' -s 2 -c 1 -f -o ${O}_l16 python tools/tune_lin.py lanes16: > ${O}_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i ${O}_l16.ncu-rep --page raw --csv > ${O}_l16_raw.csv 2>/dev/null
python tools/ncu_summary.py ${O}_l16_raw.csv 100000000 "lanes16" 2>&1 | head -70
