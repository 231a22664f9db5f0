#!/usr/bin/env bash
# round 2, 8-GPU session: bit-for-bit verifier under torchrun, the product's own multi-GPU API
# (tests + BASELINE configs, cfg4 at its full 1e9 shots), the torchrun config runner, bench at N=8
set -u
N=${1:-8}
mkdir -p gpurun_out
O=gpurun_out/r02_multi${N}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi topo -m > ${O}_topo.log 2>&1
$TR --master-port 29520 tools/verify_multi.py > ${O}_verify.log 2>&1; echo "verify rc=$?" | tee -a ${O}_verify.log
grep "^{" ${O}_verify.log | cut -c1-140
python -m pytest tests/test_gpu_multi.py -m gpu -x -q > ${O}_pytest.log 2>&1; echo "pytest multi rc=$?" | tee -a ${O}_pytest.log
tail -n 3 ${O}_pytest.log
python tools/multi_product.py > ${O}_product.log 2>&1; echo "product rc=$?" | tee -a ${O}_product.log
grep "^{" ${O}_product.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['n_gpus'], d['config'][:52], '%.3e' % (d.get('shots_per_s') or d.get('faults_per_s')), '%.3f ms' % d['ms'], d['tallies_equal_1gpu'])
"
tail -n 3 ${O}_product.log | cut -c1-300
$TR --master-port 29512 tools/run_configs.py > ${O}_configs.log 2> ${O}_configs.err; echo "configs rc=$?"
grep '^{' ${O}_configs.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['config'][:60], '%.3e' % (d.get('shots_per_s') or d.get('faults_per_s')), '%.3f ms' % d['ms'])
"
$TR --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline > ${O}_bench.log 2> ${O}_bench.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d = json.loads(open("${O}_bench.log").read().strip().splitlines()[-1])
    print("bench N=$N value %.3e  e2e %.3e (%s ms)  dense %.3e  enum %.3e (%.4f ms)" % (d["value"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e_dense"]["value"], d["enumerator"]["value"], d["enumerator"]["ms"]))
except Exception as e:
    print("bench parse failed", e)
PY
tail -c 400 ${O}_bench.err
