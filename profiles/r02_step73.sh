#!/usr/bin/env bash
# round 2, step 73: bench.py on 8 GPUs at the head of the round (16-bit events, the new creation path)
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step73_n8
timeout 45 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29543 \
  bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu-baseline --no-frame --plugin-shots 0 > ${O}_bench.log 2> ${O}_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02_step73_n8_bench.log").read().strip().splitlines()[-1])
print("n_gpus", d["n_gpus"], "value", d["value"], "ms", d["ms_per_step"])
print("e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], "dense", d["e2e_dense"]["value"])
print("enum", d["enumerator"]["value"], d["enumerator"].get("order3", {}).get("value"))
PY
