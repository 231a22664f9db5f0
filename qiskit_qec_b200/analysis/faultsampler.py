"""Quantum circuit fault path sampler (driver over ``_CFaultSampler``).

Mirror of the reference ``FaultSampler`` for ``method="propagator"`` (reference
``analysis/faultsampler.py:32-92`` constructor, ``:204-223`` sample, ``:162-202``
sample_one).  Sampling is done on the GPU with counter-based Philox, so the stream differs
from the reference's mt19937 stream; the distribution does not (tests hold it to 5 sigma)."""

from __future__ import annotations

from qiskit_qec_b200.analysis.errorpropagator import ErrorPropagator
from qiskit_qec_b200.analysis.extensions import C_FAULT_SAMPLER
from qiskit_qec_b200.analysis.faultenumerator import _with_separators
from qiskit_qec_b200.analysis.lowering import lower
from qiskit_qec_b200.locations import tag_fault_locations

if C_FAULT_SAMPLER:
    from qiskit_qec_b200.analysis.extensions import _CFaultSampler


class FaultSampler:
    """Samples faults in a circuit according to a noise model."""

    METHOD_STABILIZER: str = "stabilizer"
    METHOD_PROPAGATOR: str = "propagator"
    AVAILABLE_METHODS = {METHOD_STABILIZER, METHOD_PROPAGATOR}

    def __init__(self, circ, model, method: str = METHOD_PROPAGATOR, sim_seed: int = 0):
        self.model = model
        if method not in self.AVAILABLE_METHODS:
            raise ValueError(f"method {method} is not supported.")
        if method != self.METHOD_PROPAGATOR:
            raise NotImplementedError('only method="propagator" is accelerated; "stabilizer" is '
                                      "the reference's Aer path and is out of scope here")
        if not C_FAULT_SAMPLER:
            raise ImportError("B200 FaultSampler extension not loaded (no CPU fallback).")
        self.method = method
        self.location_types = self.model.get_operations()
        self.pauli_error_types = self.model.get_pauli_error_types()
        self.sim_seed = sim_seed
        self.circuit = lower(circ)
        self.faulty_ops_indices, self.faulty_ops_labels = tag_fault_locations(
            self.circuit.labels, self.model)
        self.label_to_pauli_weight = {
            label: {pauli: self.model.get_pauli_weight(label, pauli)
                    for pauli in self.pauli_error_types[label]}
            for label in self.faulty_ops_labels}
        self.label_to_pauli_weight_tuple = {
            label: list(pdict.items()) for label, pdict in self.label_to_pauli_weight.items()}
        self.label_to_error_probability = {
            label: self.model.get_error_probability(label) for label in self.faulty_ops_labels}
        self.propagator = ErrorPropagator()
        self.propagator.load_circuit(self.circuit)
        self.reg_sizes = list(self.circuit.creg_sizes) or [self.circuit.nc]
        self.use_compiled = True
        self.faultsamp = _CFaultSampler(
            self.circuit.nq, self.circuit.nc, self.circuit.ops, self.faulty_ops_indices,
            self.faulty_ops_labels, self.label_to_pauli_weight_tuple,
            self.label_to_error_probability, self.sim_seed)

    def sample_one(self):
        """One shot, (-1, labels, errors, outcome) with register separators as the
        reference's sample_one (``:186-202``)."""
        index, labels, errors, outcome = self.faultsamp.sample(1)[0]
        return (index, labels, errors, _with_separators(outcome, self.reg_sizes))

    def sample(self, blocksize: int = 10000):
        """[(-1, labels, errors, outcome), ...] with ``blocksize`` elements (flat outcomes)."""
        return self.faultsamp.sample(blocksize)

    # bulk API -----------------------------------------------------------------------------
    def sample_packed(self, n: int, want_bits: bool = True, want_tally: bool = False):
        """Packed outcomes uint32[nc][ceil(n/32)] and/or tallies of ``n`` fresh shots."""
        return self.faultsamp.sample_packed(n, want_bits, want_tally)
