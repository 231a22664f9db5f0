#!/usr/bin/env bash
# round 2, step 6: drop-in tests (sample() rebuilt), then the bench line and the reference arm
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step6
python -m pytest tests/test_gpu_dropin.py tests/test_gpu_configs.py -m gpu -x -q > ${O}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a ${O}_pytest.log
tail -n 6 ${O}_pytest.log
python bench.py --steps 10 --warmup 3 > ${O}_bench.log 2> ${O}_bench.err; echo "bench rc=$?"
tail -c 600 ${O}_bench.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02_step6_bench.log").read().strip().splitlines()[-1])
print("value", d["value"], "ms", d["ms_per_step"], "hbm frac", d["roofline"]["frac"])
print("e2e compact", d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e"]["d2h_bytes_per_step"])
print("e2e dense  ", d["e2e_dense"]["value"], d["e2e_dense"]["ms_per_step"])
print("plugin     ", d["e2e_plugin"])
print("enum       ", d["enumerator"]["value"], d["enumerator"]["ms"], d["enumerator"]["hbm_frac"])
for k in ("sampler", "enumerator"):
    print("frame", k, d["frame"][k]["value"], d["frame"][k]["ms"], d["frame"][k]["int_roofline"]["frac"])
print("int peak", d["int_peak"]["value"]); print("cpu", d["cpu_baseline"]); print("clocks", d["clocks"])
PY
