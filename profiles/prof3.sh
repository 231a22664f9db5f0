set -u
mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k regex:'k_sample_l' -s 3 -c 1 -f -o gpurun_out/full_lanes python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-enum --e2e-steps 1 > gpurun_out/ncu_full_lanes.log 2>&1
tail -2 gpurun_out/ncu_full_lanes.log
