"""Where a circuit + sampler creation spends its time (QEC_B200_TRACE=1 stamps on stderr):
  QEC_B200_TRACE=1 python tools/create_trace.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from qiskit_qec_b200 import _capi
circ, model, idx, labels, types = bench.make_inputs()
names = list(dict.fromkeys(labels)); cls = {n: i for i, n in enumerate(names)}
import numpy as np
# the C ABI's host buffers, laid out once (as bench.py's end-to-end legs do); "lists" = the same
# create calls fed with the Python lists of the reference's API (flattened inside the timed call)
csr = _capi.rows_to_csr(circ.ops)
loc_op = np.ascontiguousarray(idx, dtype=np.int32)
loc_cls = np.asarray([cls[l] for l in labels], dtype=np.int32)
cp = [model.get_error_probability(n) for n in names]
cc = [[_capi.pauli_code(p) for p in types[n]] for n in names]
cw = [[model.get_pauli_weight(n, p) for p in types[n]] for n in names]
for form in ("arrays", "lists"):
    for i in range(4):
        print(f"--- create #{i} ({form})", file=sys.stderr, flush=True)
        t0 = time.perf_counter()
        g = _capi.Circuit(circ.nq, circ.nc, csr if form == "arrays" else circ.ops, device=0)
        t1 = time.perf_counter()
        if form == "arrays":
            s = _capi.Sampler(g, loc_op, loc_cls, cp, cc, cw, 0)
        else:
            s = _capi.Sampler(g, idx, [cls[l] for l in labels], [model.get_error_probability(n) for n in names],
                              [[_capi.pauli_code(p) for p in types[n]] for n in names],
                              [[model.get_pauli_weight(n, p) for p in types[n]] for n in names], 0)
        t2 = time.perf_counter()
        s.close(); g.close()
        t3 = time.perf_counter()
        print(f"{form}: circuit {1e3*(t1-t0):.3f} ms  sampler {1e3*(t2-t1):.3f} ms  close {1e3*(t3-t2):.3f} ms",
              file=sys.stderr, flush=True)
