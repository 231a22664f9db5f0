set -u
mkdir -p gpurun_out
A="--no-cpu-baseline --e2e-steps 1"
python bench.py --steps 3 --warmup 3 --shots 20000000 $A > gpurun_out/p1_jit_2e7.log 2>&1
QEC_B200_JIT=0 python bench.py --steps 3 --warmup 3 --shots 20000000 $A > gpurun_out/p1_tab_2e7.log 2>&1
QEC_B200_JIT=0 python bench.py --steps 3 --warmup 3 --shots 100000000 $A > gpurun_out/p1_tab_1e8.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'qec_jit_kernel' -c 3 -f -o gpurun_out/p1_jit python bench.py --steps 1 --warmup 1 --shots 20000000 $A > gpurun_out/p1_ncu.log 2>&1
tail -c 1500 gpurun_out/p1_jit_2e7.log; echo; tail -c 1200 gpurun_out/p1_tab_2e7.log; echo; tail -c 1200 gpurun_out/p1_tab_1e8.log; tail -3 gpurun_out/p1_ncu.log
