set -u
mkdir -p gpurun_out
A="--no-cpu-baseline --no-enum --e2e-steps 1 --steps 5 --warmup 3"
run() { name=$1; shift; env "$@" python bench.py $A > gpurun_out/v_$name.log 2>&1; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/v_$name.log").read().strip().splitlines()[-1])
    print("$name", d["kernel_ms_per_launch"], d["roofline"]["kernel"])
except Exception as e: print("$name", "ERR", e)
PY
}
run base X=1
run lanes7 QEC_B200_LIN_WARPS=7
python -m pytest tests/test_gpu_linear.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
