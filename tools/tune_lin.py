#!/usr/bin/env python
"""Kernel-only sweep of the linear-response sampler variants (BASELINE configs[1]: d=5 T=5,
p=1e-3, 1e8 shots, packed outcomes + tallies on the device): ms per launch by CUDA events inside
the library, for each QEC_B200_LIN_MODE / QEC_B200_LIN_WARPS combination given on the command
line (default: a standard set).  Prints one line per combination."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import make_inputs  # noqa: E402
from qiskit_qec_b200 import _capi  # noqa: E402


def run(mode, warps, n_shots=100_000_000, tally=True, bits=True, reps=5, d=5, T=5, np_=0):
    for k in ("QEC_B200_LIN_MODE", "QEC_B200_LIN_WARPS", "QEC_B200_LANES_NP"):
        os.environ.pop(k, None)
    if np_:
        os.environ["QEC_B200_LANES_NP"] = str(np_)
    if mode:
        os.environ["QEC_B200_LIN_MODE"] = mode
    if warps:
        os.environ["QEC_B200_LIN_WARPS"] = str(warps)
    circ, model, idx, labels, types = make_inputs(d, T)
    names = list(dict.fromkeys(labels))
    cls = {n: i for i, n in enumerate(names)}
    g = _capi.Circuit(circ.nq, circ.nc, circ.ops)
    try:
        sm = _capi.Sampler(g, idx, [cls[lab] for lab in labels],
                           [model.get_error_probability(n) for n in names],
                           [[_capi.pauli_code(p) for p in types[n]] for n in names],
                           [[model.get_pauli_weight(n, p) for p in types[n]] for n in names], 0)
    except RuntimeError as e:
        print(f"mode={mode} warps={warps}: {e}")
        return
    dev = torch.device("cuda", 0)
    nw = (n_shots + 31) // 32
    dbits = torch.empty((circ.nc, nw), dtype=torch.int32, device=dev) if bits else None
    dtal = torch.zeros(_capi.tally_len(circ.nc), dtype=torch.int64, device=dev) if tally else None
    times = []
    for i in range(reps + 2):
        if tally:
            dtal.zero_()
        sm.sample_dev(i * nw * 32, n_shots, dbits.data_ptr() if bits else 0, dtal.data_ptr() if tally else 0)
        torch.cuda.synchronize()
        times.append(_capi.last_kernel_ms(0))
    ms = min(times[2:])
    print(f"mode={mode} warps={warps} np={np_} bits={int(bits)} tally={int(tally)} d={d}: {ms:.4f} ms "
          f"({n_shots / ms * 1e3:.3e} shots/s, {n_shots * circ.nc / 8 / ms / 1e6:.0f} GB/s)  {_capi.last_kernel()}",
          flush=True)
    sm.close(); g.close()


if __name__ == "__main__":
    if len(sys.argv) > 1:
        for spec in sys.argv[1:]:
            spec, _, opt = spec.partition("!")      # !nt = no tallies, !nb = no packed bits
            m, _, w = spec.partition(":")
            w, _, npx = w.partition("/")
            npx, _, dt = npx.partition("@")     # mode:warps/np@d,T
            dd, tt = (int(x) for x in dt.split(",")) if dt else (5, 5)
            run(m or None, int(w) if w else 0, np_=int(npx) if npx else 0, d=dd, T=tt,
                n_shots=100_000_000 if dd <= 5 else 20_000_000, tally=opt != "nt", bits=opt != "nb")
    else:
        run("lanes", 0)
        for w in (8, 10, 12, 14, 15, 16):
            run("lanes16", w)
        run("lanes16", 0, tally=False)
        run("lanes16", 0, bits=False)
