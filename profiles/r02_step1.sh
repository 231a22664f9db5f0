#!/usr/bin/env bash
# round 2, first GPU call: the whole GPU test-suite (new parity tests included), multi-GPU
# verifier at world size 1, the corrected integer-pipe peak, and compute-sanitizer over smoke()
# and a reduced test subset.  Outputs: gpurun_out/r02_step1_*.
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step1
python -m pytest tests -m gpu -x -q > ${O}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a ${O}_pytest.log
tail -n 25 ${O}_pytest.log
python tools/verify_multi.py > ${O}_verify1.log 2>&1; echo "verify rc=$?" | tee -a ${O}_verify1.log
cat ${O}_verify1.log | tail -8
python - > ${O}_intpeak.log 2>&1 <<'PY'
from qiskit_qec_b200 import _capi
for _ in range(3):
    print("int peak lane-ops/s, clock MHz:", _capi.measure_int_peak(0, 60.0))
PY
cat ${O}_intpeak.log
SAN=/usr/local/cuda/bin/compute-sanitizer
timeout 900 $SAN --tool memcheck --print-limit 20 python __graft_entry__.py smoke > ${O}_san_memcheck_smoke.log 2>&1; echo "memcheck smoke rc=$?"
tail -n 4 ${O}_san_memcheck_smoke.log
timeout 900 $SAN --tool racecheck --print-limit 20 python __graft_entry__.py smoke > ${O}_san_racecheck_smoke.log 2>&1; echo "racecheck smoke rc=$?"
tail -n 4 ${O}_san_racecheck_smoke.log
timeout 1500 $SAN --tool memcheck --print-limit 20 python -m pytest tests/test_gpu_linear.py -x -q -k "rep or surf3" > ${O}_san_memcheck_linear.log 2>&1; echo "memcheck linear rc=$?"
tail -n 6 ${O}_san_memcheck_linear.log
timeout 1500 $SAN --tool racecheck --print-limit 20 python -m pytest tests/test_gpu_linear.py -x -q -k "rep-0.05-9000 or surf3-0.01-20000" > ${O}_san_racecheck_linear.log 2>&1; echo "racecheck linear rc=$?"
tail -n 6 ${O}_san_racecheck_linear.log
