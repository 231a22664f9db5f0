#!/usr/bin/env bash
# round 2, step 71 (head of the round): row-major list extraction (one pass for the time-difference
# lists), SWAR counts -- the whole GPU suite, smoke(), the default bench run, the create trace
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step71
timeout 420 python -m pytest tests -m gpu -x -q > ${O}_pytest.log 2>&1; rc=$?; echo "pytest rc=$rc" | tee -a ${O}_pytest.log
tail -n 4 ${O}_pytest.log
if [ $rc -ne 0 ]; then exit $rc; fi
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > ${O}_smoke.log 2>&1; tail -n 1 ${O}_smoke.log
timeout 300 python bench.py > ${O}_bench.log 2> ${O}_bench.err; echo "bench rc=$?"
QEC_B200_TRACE=1 timeout 120 python tools/create_trace.py > ${O}_trace.log 2>&1; echo "trace rc=$?"
grep -v "qec trace" ${O}_trace.log | tail -n 8
awk '/create #3 \(arrays\)/,/arrays: circuit/' ${O}_trace.log
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02_step71_bench.log").read().strip().splitlines()[-1])
print("sampler ms", d["ms_per_step"], "frac", d["roofline"]["frac"], "traffic", d["roofline"]["traffic"])
print("enum", d["enumerator"]["ms"], d["enumerator"]["value"])
print("e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e"]["warmup_ms"], "dense", d["e2e_dense"]["value"], "plugin", d["e2e_plugin"]["value"])
print("launches", d["gpu_launches"], d["clocks"])
PY
