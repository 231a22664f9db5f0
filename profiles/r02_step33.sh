#!/usr/bin/env bash
# round 2, step 33: 24-bit events on the wire, faster creation: tests + bench end-to-end records
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step33
python -m pytest tests/test_gpu_configs.py tests/test_gpu_dropin.py tests/test_gpu_linear.py -m gpu -x -q > ${O}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a ${O}_pytest.log
tail -n 4 ${O}_pytest.log
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-frame > ${O}_bench.log 2> ${O}_bench.err; echo "bench rc=$?"; tail -3 ${O}_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02_step33_bench.log").read().strip().splitlines()[-1])
print("sampler ms", d["ms_per_step"], "frac", d["roofline"]["frac"])
print("e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e"]["d2h_bytes_per_step"], "dense", d["e2e_dense"]["value"], d["e2e_dense"]["ms_per_step"], "plugin", d["e2e_plugin"]["value"])
PY
