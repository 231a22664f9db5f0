"""Generate tests/golden/ref_vectors.json by running the UNMODIFIED reference C++
(oracle/_ref, built by oracle/build_ref.sh from /root/reference) on seeded inputs.

Run here (where /root/reference exists):  python tests/golden/make_golden.py
The fixtures are committed; the GPU box never needs /root/reference.
"""

import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from helpers import noise_inputs, random_circuit, random_fault_lists  # noqa: E402
from oracle import ref  # noqa: E402
from qiskit_qec_b200.circuits import (arc_circuit, heavy_hex_27_arc_layout,  # noqa: E402
                                      repetition_code_circuit, surface_code_circuit)
from qiskit_qec_b200.noise import standard_depolarizing_model  # noqa: E402


def pack_rows(rows):
    """list of 0/1 lists -> list of hex strings (one per row), bit c = character order"""
    return ["".join(str(int(b)) for b in r) for r in rows]


def main():
    R = ref.module()
    out = {"propagate": [], "enumerate": []}
    # random circuits x random fault lists through _CErrorPropagator.propagate
    for seed, nq, nc, n_ops, n_paths, kmax in [(101, 4, 3, 30, 64, 3), (102, 20, 11, 200, 100, 8),
                                               (103, 49, 40, 600, 70, 12)]:
        rng = np.random.default_rng(seed)
        circ = random_circuit(rng, nq, nc, n_ops)
        paths = random_fault_lists(rng, circ, n_paths, kmax)
        ep = R._CErrorPropagator(nq, nc)
        ep.load_circuit(nq, nc, circ.ops)
        outs, frames = [], []
        for icomb, errs in paths:
            outs.append(ep.propagate(icomb, errs))
            frames.append(ep.get_qubit_array())
        out["propagate"].append({"nq": nq, "nc": nc, "ops": circ.ops,
                                 "paths": [[i, e] for i, e in paths],
                                 "outcomes": pack_rows(outs), "frames": pack_rows(frames)})
    # full enumerator streams (index, labels, errors, outcome) on the BASELINE circuit families
    links, sched, color = heavy_hex_27_arc_layout()
    cases = [("rep_d3_T2", repetition_code_circuit(3, 2), 1), ("rep_d3_T2", repetition_code_circuit(3, 2), 2),
             ("surface_d3_T1", surface_code_circuit(3, 1), 1),
             ("arc_hh27_T1", arc_circuit(links, 1, "zx", sched, color), 1)]
    model = standard_depolarizing_model()
    for name, circ, order in cases:
        idx, labels, paulis = noise_inputs(circ, model)
        fe = R._CFaultEnumerator(order, circ.nq, circ.nc, circ.ops, idx, labels, paulis)
        paths = []
        while not fe.done():
            paths += fe.enumerate(1000)
        if len(paths) > 3000:  # keep the fixture small: head, tail and a stride through the middle
            keep = sorted(set(list(range(300)) + list(range(len(paths) - 300, len(paths)))
                              + list(range(0, len(paths), 97))))
        else:
            keep = list(range(len(paths)))
        out["enumerate"].append({
            "name": name, "order": order, "nq": circ.nq, "nc": circ.nc, "ops": circ.ops,
            "loc_op": idx, "labels": labels, "paulis": paulis, "total": len(paths),
            "index": keep, "errors": [paths[i][2] for i in keep],
            "path_labels": [paths[i][1] for i in keep],
            "outcomes": pack_rows([paths[i][3] for i in keep])})
    # block-boundary behaviour of enumerate(blocksize): (index, state, done, block length)
    circ = repetition_code_circuit(3, 2)
    idx, labels, paulis = noise_inputs(circ, model)
    seqs = []
    for order, bs in [(1, 15), (1, 127), (1, 40), (2, 1000), (2, 15 * 15), (1, 0)]:
        fe = R._CFaultEnumerator(order, circ.nq, circ.nc, circ.ops, idx, labels, paulis)
        seq = []
        for _ in range(200):
            if fe.done():
                break
            blk = fe.enumerate(bs)
            seq.append([fe.get_index(), fe.get_state(), fe.done(), len(blk)])
        seqs.append({"order": order, "blocksize": bs, "seq": seq})
    out["blocks"] = {"nq": circ.nq, "nc": circ.nc, "ops": circ.ops, "loc_op": idx,
                     "labels": labels, "paulis": paulis, "cases": seqs}
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_vectors.json")
    with open(path, "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
