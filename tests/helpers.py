"""Shared test scaffolding: synthetic inputs and conversions between the reference's
per-path Python tuples and the library's packed arrays."""

from __future__ import annotations

import numpy as np

from qiskit_qec_b200 import _capi
from qiskit_qec_b200.circuits import EncodedCircuit
from qiskit_qec_b200.locations import tag_fault_locations
from qiskit_qec_b200.noise import PauliNoiseModel, standard_depolarizing_model

ONE_Q = ["x", "y", "z"]
TWO_Q = ["ix", "iy", "iz", "xi", "xx", "xy", "xz", "yi", "yx", "yy", "yz", "zi", "zx", "zy", "zz"]


def random_circuit(rng: np.random.Generator, nq: int, nc: int, n_ops: int) -> EncodedCircuit:
    """Random stream over all 10 reference opcodes (cx always on distinct qubits)."""
    circ = EncodedCircuit(nq=nq, nc=nc, name="random")
    names = ["h", "s", "x", "y", "z", "cx", "id", "reset", "measure", "barrier"]
    weights = np.array([3, 2, 1, 1, 1, 6, 1, 2, 3, 0.5])
    weights = weights / weights.sum()
    for _ in range(n_ops):
        name = names[rng.choice(len(names), p=weights)]
        if name == "cx":
            if nq < 2:
                continue
            a, b = rng.choice(nq, size=2, replace=False)
            circ.cx(int(a), int(b))
        elif name == "measure":
            circ.measure(int(rng.integers(nq)), int(rng.integers(nc)))
        elif name == "barrier":
            circ.barrier()
        else:
            circ.append(name, int(rng.integers(nq)))
    return circ


def random_fault_lists(rng, circ: EncodedCircuit, n_paths: int, kmax: int, dup_prob: float = 0.05):
    """Per path: (icomb, error strings).  Pauli strings use the op's arity; occasionally a
    duplicated op index (the reference's last-wins rule) and 'i' characters."""
    paths = []
    n_ops = len(circ.ops)
    for _ in range(n_paths):
        k = int(rng.integers(0, kmax + 1))
        ops = [int(x) for x in rng.integers(0, n_ops, size=k)]
        if k >= 1 and rng.random() < dup_prob:
            ops.append(ops[0])
        errs = []
        for op in ops:
            row = circ.ops[op]
            if row[0] == 5:
                errs.append("".join(rng.choice(list("ixyz"), size=2)))
            else:
                errs.append(str(rng.choice(list("ixyz"))))
        paths.append((ops, errs))
    return paths


def fault_lists_to_csr(paths):
    off = np.zeros(len(paths) + 1, dtype=np.int64)
    ops, codes = [], []
    for i, (icomb, errs) in enumerate(paths):
        off[i + 1] = off[i] + len(icomb)
        ops += list(icomb)
        codes += [_capi.pauli_code(e) for e in errs]
    return off, np.array(ops, dtype=np.int32), np.array(codes, dtype=np.uint8)


def noise_inputs(circ: EncodedCircuit, model: PauliNoiseModel):
    """(faulty_ops_indices, faulty_ops_labels, faulty_ops_pauli_errors) as the reference
    FaultEnumerator.__init__ builds them (faultenumerator.py:110-116)."""
    idx, labels = tag_fault_locations(circ.labels, model)
    types = model.get_pauli_error_types()
    return idx, labels, [types[lab] for lab in labels]


def sampler_inputs(circ: EncodedCircuit, model: PauliNoiseModel):
    """(indices, labels, label_to_pauli_weight_tuple, label_to_error_probability) as the
    reference FaultSampler.__init__ builds them (faultsampler.py:58-73)."""
    idx, labels = tag_fault_locations(circ.labels, model)
    types = model.get_pauli_error_types()
    lw = {lab: [(p, model.get_pauli_weight(lab, p)) for p in types[lab]] for lab in labels}
    lp = {lab: model.get_error_probability(lab) for lab in labels}
    return idx, labels, lw, lp


def make_enumerator(capi, gcirc, order, idx, paulis):
    codes = [[capi.pauli_code(p) for p in plist] for plist in paulis]
    return capi.Enumerator(gcirc, order, idx, codes)


def make_sampler(capi, gcirc, idx, labels, lw, lp, seed):
    names = list(dict.fromkeys(labels))
    cls = {n: i for i, n in enumerate(names)}
    return capi.Sampler(gcirc, idx, [cls[lab] for lab in labels], [lp.get(n, 0.0) for n in names],
                        [[capi.pauli_code(p) for p, _ in lw[n]] for n in names],
                        [[w for _, w in lw[n]] for n in names], seed), names


def ref_outcome_matrix(block):
    """reference FaultPath list -> uint8[n][nc]"""
    return np.array([t[3] for t in block], dtype=np.uint8).reshape(len(block), -1)


__all__ = [name for name in dir() if not name.startswith("_")]
