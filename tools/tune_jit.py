#!/usr/bin/env python
"""Kernel-only timing sweep of the specialised sampler / enumerator over their tuning knobs
(environment variables read when the sampler / enumerator object is created).

  python tools/tune_jit.py [--shots N] [--what sampler|enum|both] KEY=v1,v2 KEY2=v1,v2 ...
"""
import itertools
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import make_inputs  # noqa: E402
from qiskit_qec_b200 import _capi  # noqa: E402


def main():
    shots, what, d, T, sweeps = 20_000_000, "both", 5, 5, {}
    args = sys.argv[1:]
    while args:
        a = args.pop(0)
        if a == "--shots":
            shots = int(args.pop(0))
        elif a == "--what":
            what = args.pop(0)
        elif a == "--d":
            d = int(args.pop(0)); T = d
        else:
            k, v = a.split("=")
            sweeps[k] = v.split(",")
    circ, model, idx, labels, types = make_inputs(d, T)
    names = list(dict.fromkeys(labels))
    cls = {n: i for i, n in enumerate(names)}
    nc = circ.nc
    dev = torch.device("cuda", 0)
    n_words = (shots + 31) // 32
    bits = torch.empty((nc, n_words), dtype=torch.int32, device=dev)
    tally = torch.zeros(_capi.tally_len(nc), dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    keys = list(sweeps)
    for combo in itertools.product(*[sweeps[k] for k in keys]):
        for k, v in zip(keys, combo):
            os.environ[k] = v
        tag = " ".join(f"{k.replace('QEC_B200_', '')}={v}" for k, v in zip(keys, combo))
        g = _capi.Circuit(circ.nq, nc, circ.ops, device=0)
        try:
            if what in ("sampler", "both"):
                sm = _capi.Sampler(g, idx, [cls[lab] for lab in labels],
                                   [model.get_error_probability(n) for n in names],
                                   [[_capi.pauli_code(p) for p in types[n]] for n in names],
                                   [[model.get_pauli_weight(n, p) for p in types[n]] for n in names], 0)
                for i in range(2):
                    sm.sample_dev(i * n_words * 32, shots, bits.data_ptr(), tally.data_ptr(), stream)
                torch.cuda.synchronize()
                sm.check()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for i in range(3):
                    sm.sample_dev((2 + i) * n_words * 32, shots, bits.data_ptr(), tally.data_ptr(), stream)
                e1.record()
                torch.cuda.synchronize()
                sm.check()
                ms = e0.elapsed_time(e1) / 3
                print(f"sampler [{tag}] {ms:8.3f} ms  {shots / ms / 1e6:9.2f} Gshots/s  {_capi.last_kernel()}", flush=True)
                sm.close()
            if what in ("enum", "both"):
                paulis = [types[lab] for lab in labels]
                en = _capi.Enumerator(g, 2, idx, [[_capi.pauli_code(p) for p in pl] for pl in paulis])
                total = en.num_paths()
                ebits = torch.empty((nc, (total + 31) // 32), dtype=torch.int32, device=dev)
                for i in range(2):
                    en.range_dev(0, total, ebits.data_ptr(), tally.data_ptr(), stream)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for i in range(3):
                    en.range_dev(0, total, ebits.data_ptr(), tally.data_ptr(), stream)
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / 3
                print(f"enum    [{tag}] {ms:8.3f} ms  {total / ms / 1e6:9.2f} Gfaults/s  {_capi.last_kernel()}", flush=True)
                en.close()
                del ebits
        except Exception as ex:  # noqa: BLE001
            print(f"[{tag}] failed: {ex}", flush=True)
        g.close()


if __name__ == "__main__":
    main()
