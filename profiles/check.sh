set -u
python -m pytest tests/test_gpu_linear.py -m gpu -x -q -k "enumerator" 2>&1 | tail -15
python - <<'PY'
import sys, json, time
sys.path.insert(0, ".")
import torch
from qiskit_qec_b200 import _capi
import tools.tune_jit as T  # noqa
PY
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1 > gpurun_out/ce.log 2>&1
python - <<'PY'
import json
d=json.loads(open("gpurun_out/ce.log").read().strip().splitlines()[-1])
print(d["enumerator"])
PY
