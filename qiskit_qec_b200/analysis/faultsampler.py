"""Quantum circuit fault path sampler (driver over ``_CFaultSampler``).

Mirror of the reference ``FaultSampler`` for ``method="propagator"`` (reference
``analysis/faultsampler.py:32-92`` constructor, ``:204-223`` sample, ``:162-202``
sample_one).  Sampling is done on the GPU with counter-based Philox, so the stream differs
from the reference's mt19937 stream; the distribution does not (tests hold it to 5 sigma)."""

from __future__ import annotations

import numpy as np

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

    def __init__(self, circ, model, method: str = METHOD_PROPAGATOR, sim_seed: int = 0, devices=None):
        """``devices`` (new; the reference is single-device): GPU indices of this box.  With more
        than one, the bulk call :meth:`sample_packed` cuts its shots across them (one process,
        results bit-identical to one GPU); the per-shot calls stay on ``devices[0]``."""
        self.model = model
        self.devices = [int(d) for d in devices] if devices is not None else None
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
        """Packed outcomes uint32[nc][ceil(n/32)] and/or tallies of ``n`` fresh shots (on every
        device of ``devices`` when several were given)."""
        if self.devices is not None and len(self.devices) > 1:
            multi = self._multi_sampler()
            lo = self.faultsamp.get_cursor()
            out = multi.sample(lo, int(n), want_bits, want_tally)
            self.faultsamp.set_cursor(lo + (int(n) + 31) // 32 * 32)
            return out
        return self.faultsamp.sample_packed(n, want_bits, want_tally)

    def _multi_sampler(self):
        if getattr(self, "_multi", None) is None:
            from qiskit_qec_b200 import _capi
            names = list(dict.fromkeys(self.faulty_ops_labels))
            cls = {name: i for i, name in enumerate(names)}
            lw = self.label_to_pauli_weight_tuple
            self._multi = _capi.MultiSampler(
                self.devices, self.circuit.nq, self.circuit.nc, self.circuit.ops, self.faulty_ops_indices,
                [cls[lab] for lab in self.faulty_ops_labels],
                [self.label_to_error_probability.get(name, 0.0) for name in names],
                [[_capi.pauli_code(p) for p, _ in lw[name]] for name in names],
                [[w for _, w in lw[name]] for name in names], self.sim_seed)
        return self._multi

    def sample_events(self, n: int, want_tally: bool = False):
        """Compact form of ``n`` fresh shots: ``(word_cnt, events, tally)`` -- per 32-shot word its
        number of faults, and all faults back to back (``location << 14 | shot_bit << 9 | digit``,
        ``location`` indexing ``faulty_ops_indices`` / ``faulty_ops_labels``, ``digit`` the Pauli of
        that label's list).  Six times fewer bytes than the packed outcomes at p = 1e-3, and
        lossless: labels and errors are the list itself, :meth:`events_to_paths` expands it."""
        return self.faultsamp.sample_events(int(n), want_tally)

    def events_to_paths(self, word_cnt, events, n: int):
        """[(labels, errors)] per shot from :meth:`sample_events` output (host-side formatting of
        what the reference's ``sample()`` returns as its 2nd and 3rd tuple entries)."""
        out = [([], []) for _ in range(n)]
        off = 0
        for w, k in enumerate(np.asarray(word_cnt).tolist()):
            for e in np.sort(np.asarray(events[off:off + k])).tolist():
                loc, bit, dig = e >> 14, (e >> 9) & 31, e & 511
                shot = 32 * w + bit
                if shot < n:
                    label = self.faulty_ops_labels[loc]
                    out[shot][0].append(label)
                    out[shot][1].append(self.label_to_pauli_weight_tuple[label][dig][0])
            off += k
        return out

    def sample_counts(self, shots: int = 4000):
        """Counts dictionary in standard Qiskit form for ``shots`` fresh shots: keys are the
        outcome bitstrings with the LAST clbit first and a space between classical registers,
        values their multiplicities -- the format of the reference's
        ``get_counts_via_stim`` (``utils/stim_tools.py:234-283``) and what every decoder of the
        reference consumes (``decoders/decoding_graph.py:165-209``).  Outcomes are FLIPS
        relative to the noiseless run (the propagator's semantics): they equal the stim /
        Aer absolute outcomes when the noiseless outcome is all zero (base circuit "0").
        Sampling AND counting run on the GPU (K4, the hashed outcome table of
        ``qec_sampler_counts``): only the distinct outcomes cross PCIe; the host formats strings."""
        if shots <= 0:
            return {}
        rows, counts, n_zero = self.faultsamp.sample_counts(int(shots))
        return counts_from_rows(np.asarray(rows), np.asarray(counts), int(n_zero),
                                self.circuit.nc, self.reg_sizes)


def _format_key(chars, cuts) -> str:
    return "".join((" " if i in cuts else "") + ("1" if b else "0") for i, b in enumerate(chars))


def counts_from_rows(rows: np.ndarray, counts: np.ndarray, n_zero: int, nc: int, reg_sizes) -> dict:
    """{bitstring: count} from K4's table: ``rows`` uint32[k][ceil(nc/32)] (bit c % 32 of word
    c / 32 = clbit c) with multiplicities ``counts``, plus ``n_zero`` all-zero outcomes.  Strings
    list clbit nc-1 first, registers (declaration order ``reg_sizes``) separated by a space."""
    if sum(reg_sizes) != nc:
        raise ValueError("register sizes do not add up to the number of clbits")
    cuts = set(np.cumsum(list(reg_sizes)[::-1])[:-1].tolist())  # after these many chars: a space
    out = {}
    if n_zero:
        out[_format_key(np.zeros(nc, dtype=np.uint8), cuts)] = int(n_zero)
    if len(rows):
        bits = np.unpackbits(np.ascontiguousarray(rows).view(np.uint8), axis=1, bitorder="little")[:, :nc]
        for chars, k in zip(bits[:, ::-1], counts):
            s = _format_key(chars, cuts)
            out[s] = out.get(s, 0) + int(k)
    return out


def counts_from_packed(packed: np.ndarray, n: int, reg_sizes) -> dict:
    """{bitstring: count} from packed outcomes uint32[nc][ceil(n/32)] (bit b of word w of row c
    = clbit c of shot 32 w + b), counted on the host with numpy: the CHECKER of
    :func:`counts_from_rows` in the tests and a convenience for outcome buffers that are already
    on the host."""
    nc = packed.shape[0]
    if sum(reg_sizes) != nc:
        raise ValueError("register sizes do not add up to the number of clbits")
    bits = np.unpackbits(np.ascontiguousarray(packed).view(np.uint8).reshape(nc, -1), axis=1,
                         bitorder="little")[:, :n]            # [clbit][shot]
    rows = np.packbits(bits[::-1].T, axis=1)                   # per shot, clbit nc-1 first
    uniq, cnt = np.unique(rows, axis=0, return_counts=True)
    cuts = set(np.cumsum(list(reg_sizes)[::-1])[:-1].tolist())
    return {_format_key(np.unpackbits(u)[:nc], cuts): int(k) for u, k in zip(uniq, cnt)}
