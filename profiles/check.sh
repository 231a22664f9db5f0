set -u
python -m pytest tests/test_gpu_linear.py -m gpu -x -q -k "enumerator" 2>&1 | tail -3
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1 > gpurun_out/ce.log 2>&1
python - <<'PY'
import json
d=json.loads(open("gpurun_out/ce.log").read().strip().splitlines()[-1])
print(d["enumerator"]); print(d["kernel_ms_per_launch"])
PY
