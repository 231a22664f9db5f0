set -u
N=$1
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${N}gpu.log 2> gpurun_out/bench_${N}gpu.err
tail -c 900 gpurun_out/bench_${N}gpu.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/run_configs.py > gpurun_out/configs_${N}gpu.log 2> gpurun_out/configs_${N}gpu.err
grep '^{' gpurun_out/configs_${N}gpu.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['config'][:40], d.get('shots_per_s') or d.get('faults_per_s'), d['ms'])
"
