"""minimum_distance on the rotated surface code d=5 (25 qubits): the GPU brute force against the
compiled reference (oracle/_ref), same inputs.  The reference is timed at max_weight 3 and 4 only
(weight 5 alone is 1.3e7 errors with one rank computation each)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from qiskit_qec_b200.analysis import _c_analysis as ours
from oracle import ref as _ref
ref = _ref.module() if _ref.available() else None


def surface(d):
    q = lambda r, c: r * d + c
    kind = lambda r, c: "x" if (r + c) % 2 == 0 else "z"
    st = [(kind(r, c), [q(r, c), q(r, c + 1), q(r + 1, c), q(r + 1, c + 1)]) for r in range(d - 1) for c in range(d - 1)]
    for c in range(d - 1):
        if kind(0, c) == "z": st.append(("x", [q(0, c), q(0, c + 1)]))
        if kind(d - 2, c) == "z": st.append(("x", [q(d - 1, c), q(d - 1, c + 1)]))
    for r in range(d - 1):
        if kind(r, 0) == "x": st.append(("z", [q(r, 0), q(r + 1, 0)]))
        if kind(r, d - 2) == "x": st.append(("z", [q(r, d - 1), q(r + 1, d - 1)]))
    n = d * d
    m = np.zeros((len(st), 2 * n), dtype=int)
    for i, (k, qs) in enumerate(st):
        for x in qs:
            m[i, x + (0 if k == "x" else n)] = 1
    return m.tolist()


for d, weights in ((5, (3, 4, 6)), (7, (3, 4, 5))):
    stab = surface(d)
    ours._c_minimum_distance(stab, stab, 1)  # warm-up (context, tables)
    for mw in weights:
        t0 = time.perf_counter(); got = ours._c_minimum_distance(stab, stab, mw); t1 = time.perf_counter()
        line = f"d={d} max_weight={mw}: GPU {got} in {1e3 * (t1 - t0):.1f} ms"
        if ref is not None and ((d == 5 and mw <= 4) or (d == 7 and mw <= 3)):
            t0 = time.perf_counter(); want = ref._c_minimum_distance(stab, stab, mw); t1 = time.perf_counter()
            line += f"   reference {want} in {t1 - t0:.2f} s"
        print(line, flush=True)
