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
run lanes4 QEC_B200_LIN_WARPS=4
run warp QEC_B200_LIN_MODE=warp
run lanes6 QEC_B200_LIN_WARPS=6

ncu --set full --clock-control none --import-source on -k regex:'k_sample_l' -s 3 -c 1 -f -o gpurun_out/full_lean python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-enum --e2e-steps 1 > gpurun_out/ncu_full_lean.log 2>&1
tail -2 gpurun_out/ncu_full_lean.log
