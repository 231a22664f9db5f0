#!/usr/bin/env python
"""The BASELINE multi-GPU configurations through the PRODUCT's own multi-GPU API (one process,
qec_multi_*: contiguous shards, tallies reduced on device 0 over peer mappings) -- not through
torchrun.  For each device count: device-timed rates (events on device 0 around every device's
kernels + the reduce) and a check that the tallies equal the 1-device tallies entry by entry.

  python tools/multi_product.py [--shots4 1000000000]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from qiskit_qec_b200 import _capi  # noqa: E402
from qiskit_qec_b200.circuits import arc_circuit, heavy_hex_27_arc_layout, surface_code_circuit  # noqa: E402
from qiskit_qec_b200.locations import tag_fault_locations  # noqa: E402
from qiskit_qec_b200.noise import standard_depolarizing_model  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shots2", type=int, default=800_000_000)
    ap.add_argument("--shots4", type=int, default=1_000_000_000)
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    n_all = _capi.device_count()
    counts = [k for k in (1, 2, 4, 8) if k <= n_all]

    def sampler(name, circ, n_shots, p=1e-3):
        model = standard_depolarizing_model(p)
        idx, labels = tag_fault_locations(circ.labels, model)
        types = model.get_pauli_error_types()
        names = list(dict.fromkeys(labels))
        cls = {n: i for i, n in enumerate(names)}
        base = None
        for k in counts:
            ms = _capi.MultiSampler(list(range(k)), circ.nq, circ.nc, circ.ops, idx, [cls[lab] for lab in labels],
                                    [model.get_error_probability(n) for n in names],
                                    [[_capi.pauli_code(q) for q in types[n]] for n in names],
                                    [[model.get_pauli_weight(n, q) for q in types[n]] for n in names], 7)
            ms.tally_dev(0, min(n_shots, 10_000_000))
            times = []
            for _ in range(args.reps):
                tally, t = ms.tally_dev(0, n_shots)
                times.append(t)
            assert int(tally[circ.nc + 1]) == n_shots
            base = tally if base is None else base
            t = float(np.median(times))
            print(json.dumps({"config": name, "api": "qec_multi_sampler_tally_dev", "n_gpus": k, "shots": n_shots,
                              "ms": t, "shots_per_s": n_shots / t * 1e3, "tallies_equal_1gpu": bool(np.array_equal(tally, base)),
                              "mean_faults_per_shot": float((tally[circ.nc + 2:] * np.arange(64)).sum() / n_shots),
                              "nonzero_outcomes": int(tally[circ.nc]), "kernel": _capi.last_kernel()}), flush=True)
            ms.close()

    def enumerator(name, circ, order, limit=None):
        model = standard_depolarizing_model()
        idx, labels = tag_fault_locations(circ.labels, model)
        types = model.get_pauli_error_types()
        codes = [[_capi.pauli_code(q) for q in types[lab]] for lab in labels]
        base = None
        for k in counts:
            me = _capi.MultiEnumerator(list(range(k)), circ.nq, circ.nc, circ.ops, order, idx, codes)
            total = me.num_paths()
            n = min(total, limit) if limit else total
            me.tally_dev(0, min(n, 1_000_000))
            times = []
            for _ in range(args.reps):
                tally, t = me.tally_dev(0, n)
                times.append(t)
            assert int(tally[circ.nc + 1]) == n
            base = tally if base is None else base
            t = float(np.median(times))
            print(json.dumps({"config": name, "api": "qec_multi_enumerator_tally_dev", "n_gpus": k, "paths_total": total,
                              "paths_run": n, "ms": t, "faults_per_s": n / t * 1e3,
                              "tallies_equal_1gpu": bool(np.array_equal(tally, base)), "kernel": _capi.last_kernel()}),
                  flush=True)
            me.close()

    sampler("cfg2 surface d=5 T=5 sampler, tallies", surface_code_circuit(5, 5), args.shots2)
    enumerator("cfg3 surface d=5 T=5 order 2 (21.7M paths), tallies", surface_code_circuit(5, 5), 2)
    enumerator("cfg3+ surface d=5 T=5 order 3, first 4e9 of 4.8e10 paths, tallies", surface_code_circuit(5, 5), 3,
               limit=4_000_000_000)
    sampler(f"cfg4 surface d=11 T=11 sampler, FULL {args.shots4} shots, tallies", surface_code_circuit(11, 11), args.shots4)
    links, schedule, color = heavy_hex_27_arc_layout()
    enumerator("cfg5 heavy-hex ARC T=3 order 3, ALL paths, tallies", arc_circuit(links, 3, "zx", schedule, color), 3)


if __name__ == "__main__":
    main()
