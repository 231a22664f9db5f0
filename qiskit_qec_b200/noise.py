"""Circuit-level Pauli noise model (host side of the hot path's inputs).

Mirror of the reference ``PauliNoiseModel`` interface
(reference ``src/qiskit_qec/noise/paulinoisemodel.py:8-141``): same method names,
argument meaning, normalisation on insert and error behaviour.  The reference's only
qiskit_aer use (``as_aer_noise_model``, ``:143-163``) is outside the hot path and omitted.
"""

from __future__ import annotations

import copy
from typing import Dict, List, Union

_PAULI_ALPHABET = set("ixyz")


class PauliNoiseModel:
    """``{label: {pauli_string: weight}}`` + per-label error probability and scale factor."""

    def __init__(self, fromdict=None):
        self.definition: Dict[str, Dict[str, float]] = {}
        self.error_probabilities: Dict[str, float] = {}
        self.scale_factors: Dict[str, float] = {}
        self.num_qubits: Dict[str, int] = {}
        self.chan_label = "chan"
        self.p_label = "p"
        self.scale_label = "scale"
        if fromdict is not None:
            for name, propdict in fromdict.items():
                assert self.chan_label in propdict
                self.add_operation(name, propdict[self.chan_label])
                if propdict.get(self.p_label) is not None:
                    self.set_error_probability(name, propdict[self.p_label])
                if propdict.get(self.scale_label) is not None:
                    self.set_scale_factor(name, propdict[self.scale_label])

    def to_dict(self):
        """Inverse of ``__init__(fromdict=...)`` (reference ``:32-43``)."""
        out = {}
        for name, channel in self.definition.items():
            out[name] = {self.chan_label: copy.deepcopy(channel)}
            if name in self.error_probabilities:
                out[name][self.p_label] = self.error_probabilities[name]
            if name in self.scale_factors:
                out[name][self.scale_label] = self.scale_factors[name]
        return out

    def get_operations(self) -> List[str]:
        """Labels of faulty operations, insertion ordered (reference ``:45-47``)."""
        return list(self.definition.keys())

    def get_pauli_error_types(self) -> Dict[str, List[str]]:
        """Per label, the Pauli strings in insertion order (reference ``:49-54``).

        This order defines the digit order of the enumerator's per-location product.
        """
        return {k: list(v.keys()) for k, v in self.definition.items()}

    def get_error_probability(self, name: str) -> float:
        if name not in self.error_probabilities:
            raise Exception(f'no error probability for "{name}"')
        return self.error_probabilities[name]

    def get_pauli_weight(self, name: str, paulistring: str) -> float:
        if name not in self.definition:
            raise Exception(f'"{name}" is not an operation')
        if set(paulistring) > _PAULI_ALPHABET:
            raise Exception(f'bad paulistring "{paulistring}"')
        if len(paulistring) != self.num_qubits[name]:
            raise Exception(f"paulistring not on {self.num_qubits[name]} qubits")
        return self.definition[name].get(paulistring, 0)

    def get_pauli_error_probability(self, name: str, paulistring: str) -> float:
        return self.get_error_probability(name) * self.get_pauli_weight(name, paulistring)

    def set_scale_factor(self, name: str, factor: Union[int, float]) -> None:
        factor = float(factor)
        if name not in self.definition:
            raise Exception(f'"{name}" is not an operation')
        if factor < 0:
            raise Exception("expected non-negative factor")
        self.scale_factors[name] = factor

    def set_scaled_error_probabilities(self, p: float) -> None:
        for name in self.get_operations():
            if name in self.scale_factors:
                self.set_error_probability(name, self.scale_factors[name] * p)

    def set_error_probability(self, name: str, p: float) -> None:
        if name not in self.definition:
            raise Exception(f'"{name}" is not an operation')
        if p < 0 or p > 1:
            raise Exception("expected a probability")
        self.error_probabilities[name] = p

    def add_operation(self, name: str, paulichanneldict: Dict[str, float]) -> None:
        """Add a faulty operation; weights are normalised on insert (reference ``:114-141``)."""
        if len(paulichanneldict) == 0:
            raise Exception("expected non-empty dictionary")
        total = 0
        nq = -1
        for k, v in paulichanneldict.items():
            if set(k) > _PAULI_ALPHABET:
                raise Exception(f'bad paulistring "{k}"')
            if nq == -1:
                nq = len(k)
            elif len(k) != nq:
                raise Exception(f"expected paulistring on {nq} qubits")
            total += v
        self.definition[name] = {k: v / total for k, v in paulichanneldict.items()}
        self.num_qubits[name] = nq


TWO_QUBIT_PAULIS = ["ix", "iy", "iz", "xi", "xx", "xy", "xz", "yi", "yx", "yy", "yz",
                    "zi", "zx", "zy", "zz"]


def standard_depolarizing_model(p: float = None, labels=("h", "s", "x", "y", "z", "id")) -> PauliNoiseModel:
    """The default model of the reference's ``FaultEnumerator`` (``faultenumerator.py:64-93``):
    xyz on the listed 1-qubit gates, x on reset and measure, the 15 two-qubit Paulis on cx.
    If ``p`` is given it is assigned to every label (circuit-level noise of SURVEY.md 8(d))."""
    m = PauliNoiseModel()
    for name in labels:
        m.add_operation(name, {"x": 1, "y": 1, "z": 1})
    m.add_operation("reset", {"x": 1})
    m.add_operation("measure", {"x": 1})
    m.add_operation("cx", {k: 1 for k in TWO_QUBIT_PAULIS})
    if p is not None:
        for name in m.get_operations():
            m.set_error_probability(name, p)
    return m
