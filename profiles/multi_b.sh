set -u
N=$1
for k in 1 2; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$k bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline --no-enum > gpurun_out/bench_${N}gpu_$k.log 2> gpurun_out/bench_${N}gpu_$k.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_${N}gpu_$k.log").read().strip().splitlines()[-1])
print("N=$N run $k value", d["value"], "e2e", d["e2e"]["value"], "ms", d["ms_per_step"])
PY
done
nvidia-smi topo -m 2>/dev/null | head -12
