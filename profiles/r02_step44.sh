#!/usr/bin/env bash
# round 2, step 44: bench with the order-3 enumerator record, 1 GPU and 2 GPUs (torchrun)
set -u
N=${1:-1}
mkdir -p gpurun_out
O=gpurun_out/r02_step44_n${N}
if [ "$N" = "1" ]; then
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-frame --plugin-shots 0 > ${O}_bench.log 2> ${O}_bench.err; echo "bench rc=$?"
else
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline --no-frame --plugin-shots 0 > ${O}_bench.log 2> ${O}_bench.err; echo "bench rc=$?"
fi
tail -3 ${O}_bench.err | cut -c1-300
python - <<PY
import json
d=json.loads(open("${O}_bench.log").read().strip().splitlines()[-1])
print("N", d["n_gpus"], "sampler", "%.3e"%d["value"], d["ms_per_step"], "e2e %.3e"%d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e"]["warmup_ms"])
print("enum %.3e %.4f ms"%(d["enumerator"]["value"], d["enumerator"]["ms"]), "order3 %.3e %.3f ms"%(d["enumerator"]["order3"]["value"], d["enumerator"]["order3"]["ms"]))
PY
