#!/usr/bin/env bash
# round 2, step 18: bench line after the lane read-out / CTA read-out work (+ reference arm)
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step18
python bench.py > ${O}_bench.log 2> ${O}_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > ${O}_bench_ref.log 2>> ${O}_bench.err; echo "ref rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02_step18_bench.log").read().strip().splitlines()[-1])
print("sampler ms", d["ms_per_step"], "frac", d["roofline"]["frac"], "traffic", d["roofline"]["traffic"])
print("enum", d["enumerator"]["ms"], d["enumerator"]["hbm_frac"], d["enumerator"]["value"])
print("e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], "dense", d["e2e_dense"]["value"], "plugin", d["e2e_plugin"]["value"])
print("frame", d["frame"]["sampler"]["value"], d["frame"]["sampler"]["int_roofline"]["frac"], d["frame"]["enumerator"]["value"], d["frame"]["enumerator"]["int_roofline"]["frac"])
print("cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["one_core"]["value"])
PY
tail -c 600 ${O}_bench_ref.log
