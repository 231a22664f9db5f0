#!/usr/bin/env bash
# round 2, final single-GPU session: the whole GPU suite (-x), smoke(), the default bench run and
# the reference arm, then (after those exited) the launch list and one full capture of the
# dominant kernel for profiles/r02/ncu_dominant_kernel.json
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_final
python -m pytest tests -m gpu -x -q > ${O}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a ${O}_pytest.log
tail -n 4 ${O}_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > ${O}_smoke.log 2>&1; tail -n 1 ${O}_smoke.log
python bench.py > ${O}_bench.log 2> ${O}_bench.err; echo "bench rc=$?"
python bench.py --impl reference > ${O}_bench_ref.log 2>> ${O}_bench.err; echo "ref rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file ${O}_launches.csv \
    python bench.py --steps 5 --warmup 3 --no-cpu-baseline > ${O}_ncu_launches.log 2>&1; echo "launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_sample_lanes' -s 3 -c 1 \
    -f -o ${O}_full_lanes python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-enum --no-frame --plugin-shots 0 > ${O}_ncu_full.log 2>&1; echo "full rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02_final_bench.log").read().strip().splitlines()[-1])
print("sampler ms", d["ms_per_step"], "frac", d["roofline"]["frac"], "traffic", d["roofline"]["traffic"])
print("enum", d["enumerator"]["ms"], d["enumerator"]["hbm_frac"], d["enumerator"]["value"])
print("e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], "dense", d["e2e_dense"]["value"], "plugin", d["e2e_plugin"]["value"])
print("cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["one_core"]["value"], "launches", d["gpu_launches"], d["clocks"])
r=json.loads(open("gpurun_out/r02_final_bench_ref.log").read().strip().splitlines()[-1])
print("ref", r["value"], r["cpu_baseline"]["cores"])
PY
