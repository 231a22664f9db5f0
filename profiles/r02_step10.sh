#!/usr/bin/env bash
# round 2, step 10: lane read-out with the chain's running XOR in registers + red.shared tallies:
# whole GPU suite, then the kernel sweep and a short bench.
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step10
python -m pytest tests -m gpu -x -q > ${O}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a ${O}_pytest.log
tail -n 12 ${O}_pytest.log
python tools/tune_lin.py lanes: 'lanes:!nt' 'lanes:!nb' lanes:8 > ${O}_tune.log 2>&1; echo "tune rc=$?"
cat ${O}_tune.log
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-frame --plugin-shots 0 > ${O}_bench.log 2> ${O}_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02_step10_bench.log").read().strip().splitlines()[-1])
print("sampler ms", d["ms_per_step"], "frac", d["roofline"]["frac"], "enum", d["enumerator"]["ms"], d["enumerator"]["hbm_frac"], "e2e", d["e2e"]["value"])
PY
