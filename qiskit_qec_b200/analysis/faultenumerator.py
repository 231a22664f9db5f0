"""Quantum circuit fault path enumerator (driver over ``_CFaultEnumerator``).

Mirror of the reference ``FaultEnumerator`` for ``method="propagator"`` (reference
``analysis/faultenumerator.py:42-126`` constructor, ``:196-236`` generate, ``:238-264``
generate_blocks).  ``generate()`` is served from GPU-filled blocks instead of one
propagator call per fault; element order, indices, labels, errors, outcomes and the
per-register ``" "`` separators are the reference's."""

from __future__ import annotations

from typing import List

from qiskit_qec_b200.analysis.errorpropagator import ErrorPropagator
from qiskit_qec_b200.analysis.extensions import C_FAULT_ENUMERATOR
from qiskit_qec_b200.analysis.lowering import lower
from qiskit_qec_b200.locations import tag_fault_locations
from qiskit_qec_b200.noise import standard_depolarizing_model

if C_FAULT_ENUMERATOR:
    from qiskit_qec_b200.analysis.extensions import _CFaultEnumerator


def _with_separators(raw: List[int], reg_sizes: List[int]):
    """reference ``faultenumerator.py:224-234``: a " " between classical registers"""
    if len(reg_sizes) <= 1:
        return raw
    out, j = [], 0
    for size in reg_sizes:
        out += raw[j:j + size]
        j += size
        out.append(" ")
    out.pop(-1)
    return out


class FaultEnumerator:
    """Enumerates faults in a circuit according to a noise model."""

    METHOD_STABILIZER: str = "stabilizer"
    METHOD_PROPAGATOR: str = "propagator"
    AVAILABLE_METHODS = {METHOD_STABILIZER, METHOD_PROPAGATOR}

    def __init__(self, circ, order: int = 1, method: str = METHOD_PROPAGATOR, model=None,
                 sim_seed: int = 0, devices=None):
        """``devices`` (new; the reference is single-device): with more than one GPU index,
        :meth:`enumerate_packed` cuts its path range across them in one process."""
        self.devices = [int(d) for d in devices] if devices is not None else None
        if order < 1:
            raise Exception("order < 1")
        self.order = order
        self.model = standard_depolarizing_model() if model is None else model
        if method not in self.AVAILABLE_METHODS:
            raise ValueError(f"method {method} is not supported.")
        if method != self.METHOD_PROPAGATOR:
            raise NotImplementedError('only method="propagator" is accelerated; "stabilizer" is '
                                      "the reference's Aer path and is out of scope here")
        if not C_FAULT_ENUMERATOR:
            raise ImportError("B200 FaultEnumerator extension not loaded (no CPU fallback).")
        self.method = method
        self.location_types = self.model.get_operations()
        self.pauli_error_types = self.model.get_pauli_error_types()
        self.sim_seed = sim_seed
        self.circuit = lower(circ)
        self.faulty_ops_indices, self.faulty_ops_labels = tag_fault_locations(
            self.circuit.labels, self.model)
        self.faulty_ops_pauli_errors = [self.pauli_error_types[x] for x in self.faulty_ops_labels]
        self.propagator = ErrorPropagator()
        self.propagator.load_circuit(self.circuit)
        self.reg_sizes = list(self.circuit.creg_sizes) or [self.circuit.nc]
        self.use_compiled = True
        self.faultenum = _CFaultEnumerator(
            self.order, self.circuit.nq, self.circuit.nc, self.circuit.ops,
            self.faulty_ops_indices, self.faulty_ops_labels, self.faulty_ops_pauli_errors)

    def generate(self, blocksize: int = 1 << 16):
        """Yields (index, labels, error, outcome) per fault path, in the reference's order."""
        fe = _CFaultEnumerator(self.order, self.circuit.nq, self.circuit.nc, self.circuit.ops,
                               self.faulty_ops_indices, self.faulty_ops_labels,
                               self.faulty_ops_pauli_errors)
        while not fe.done():
            for index, labels, error, outcome in fe.enumerate(blocksize):
                yield (index, labels, error, _with_separators(outcome, self.reg_sizes))

    def generate_blocks(self, blocksize: int = 10000):
        """Yields [(index, labels, error, outcome), ...] blocks exactly as the reference's
        compiled path cuts them (flat outcomes, trailing empty block when the stream ends on a
        block boundary)."""
        while not self.faultenum.done():
            yield self.faultenum.enumerate(blocksize)

    # bulk API -----------------------------------------------------------------------------
    def num_paths(self) -> int:
        return self.faultenum.num_paths()

    def enumerate_packed(self, lo: int = 0, hi: int = None, want_tally: bool = False):
        """Packed outcomes uint32[nc][ceil((hi-lo)/32)] (+ tallies) of paths [lo, hi)."""
        hi = self.num_paths() if hi is None else hi
        if self.devices is not None and len(self.devices) > 1:
            if getattr(self, "_multi", None) is None:
                from qiskit_qec_b200 import _capi
                self._multi = _capi.MultiEnumerator(
                    self.devices, self.circuit.nq, self.circuit.nc, self.circuit.ops, self.order,
                    self.faulty_ops_indices,
                    [[_capi.pauli_code(p) for p in pl] for pl in self.faulty_ops_pauli_errors])
            bits, tally = self._multi.range(lo, hi, True, want_tally)
            return bits, tally
        return self.faultenum.enumerate_packed(lo, hi, want_tally)
