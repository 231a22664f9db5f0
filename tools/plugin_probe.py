"""Where the literal drop-in call _CFaultSampler.sample(N) spends its time: successive calls (the
first one sizes the event rows), against the packed bulk call of the same module."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from qiskit_qec_b200.analysis import _c_analysis as ext
circ, model, idx, labels, types = bench.make_inputs()
names = list(dict.fromkeys(labels))
lw = {n: [(p, float(model.get_pauli_weight(n, p))) for p in types[n]] for n in names}
lp = {n: float(model.get_error_probability(n)) for n in names}
fs = ext._CFaultSampler(circ.nq, circ.nc, circ.ops, idx, labels, lw, lp, 1)
for n in (100_000, 1_000_000, 1_000_000, 1_000_000):
    t0 = time.perf_counter(); blk = fs.sample(n); t1 = time.perf_counter()
    print(f"sample({n}): {1e3*(t1-t0):.1f} ms  {n/(t1-t0):.3e} shots/s", flush=True)
    t0 = time.perf_counter(); del blk; t1 = time.perf_counter()
    print(f"  del block: {1e3*(t1-t0):.1f} ms", flush=True)
for n in (1_000_000, 1_000_000):
    t0 = time.perf_counter(); out = fs.sample_packed(n, True, True); t1 = time.perf_counter()
    print(f"sample_packed({n}): {1e3*(t1-t0):.1f} ms", flush=True)
