#!/usr/bin/env bash
# round 2, step 23: whole-batch read-out with unaligned rows (enumerator: n_words not a multiple of 4)
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step23
python -m pytest tests/test_gpu_linear.py tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_configs.py tests/test_gpu_dropin.py -m gpu -x -q > ${O}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a ${O}_pytest.log
tail -n 4 ${O}_pytest.log
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-frame --plugin-shots 0 > ${O}_bench.log 2> ${O}_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02_step23_bench.log").read().strip().splitlines()[-1])
print("sampler ms", d["ms_per_step"], "frac", d["roofline"]["frac"], "traffic", d["roofline"]["traffic"])
print("enum", d["enumerator"]["ms"], d["enumerator"]["hbm_frac"], d["enumerator"]["value"])
PY
