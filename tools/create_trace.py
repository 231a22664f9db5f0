"""Where a circuit + sampler creation spends its time (QEC_B200_TRACE=1 stamps on stderr):
  QEC_B200_TRACE=1 python tools/create_trace.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from qiskit_qec_b200 import _capi
circ, model, idx, labels, types = bench.make_inputs()
names = list(dict.fromkeys(labels)); cls = {n: i for i, n in enumerate(names)}
for i in range(4):
    print(f"--- create #{i}", file=sys.stderr, flush=True)
    t0 = time.perf_counter()
    g = _capi.Circuit(circ.nq, circ.nc, circ.ops, device=0)
    t1 = time.perf_counter()
    s = _capi.Sampler(g, idx, [cls[l] for l in labels], [model.get_error_probability(n) for n in names],
                      [[_capi.pauli_code(p) for p in types[n]] for n in names],
                      [[model.get_pauli_weight(n, p) for p in types[n]] for n in names], 0)
    t2 = time.perf_counter()
    s.close(); g.close()
    t3 = time.perf_counter()
    print(f"circuit {1e3*(t1-t0):.2f} ms  sampler {1e3*(t2-t1):.2f} ms  close {1e3*(t3-t2):.2f} ms", file=sys.stderr, flush=True)
