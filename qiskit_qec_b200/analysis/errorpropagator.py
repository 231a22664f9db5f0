"""ErrorPropagator interface over the B200 extension.

Mirror of the reference's ``BaseErrorPropagator`` / ``CErrorPropagator`` / ``ErrorPropagator``
(reference ``analysis/baseerrorpropagator.py:7-90``, ``cerrorpropagator.py:16-130``,
``errorpropagator.py:37-62``): same method names, argument meaning and error behaviour."""

from __future__ import annotations

from typing import List, Sequence, Tuple

from qiskit_qec_b200.analysis.extensions import C_ERROR_PROPAGATOR
from qiskit_qec_b200.analysis.lowering import lower
from qiskit_qec_b200.circuits import OP_NAMES

if C_ERROR_PROPAGATOR:
    from qiskit_qec_b200.analysis.extensions import _CErrorPropagator


class CErrorPropagator:
    """ErrorPropagator object using the compiled (CUDA) extension."""

    def __init__(self, qreg_size: int = 1, creg_size: int = 1):
        if not C_ERROR_PROPAGATOR:
            raise ImportError("B200 ErrorPropagator extension not loaded (no CPU fallback).")
        self.stabilizer_op_names = list(OP_NAMES)
        self.stabilizer_op_codes = list(range(10))
        self.name_to_code = dict(zip(self.stabilizer_op_names, self.stabilizer_op_codes))
        self.encoded_circ = None
        self.circuit = None
        self.cep = _CErrorPropagator(qreg_size, creg_size)

    def apply_error(self, q_idx: List[int], err_str: str):
        assert len(q_idx) == len(err_str), "length mismatch"
        assert set(err_str) <= set("ixyz"), "bad error string"
        self.cep.apply_error(q_idx, err_str)

    def load_circuit(self, circ):
        """``circ``: qiskit ``QuantumCircuit`` or an ``EncodedCircuit``."""
        self.circuit = lower(circ)
        self.encoded_circ = self.circuit.ops
        self.cep.load_circuit(self.circuit.nq, self.circuit.nc, self.encoded_circ)

    def cx(self, qc: int, qt: int):
        self.cep.cx(qc, qt)

    def h(self, q: int):
        self.cep.h(q)

    def s(self, q: int):
        self.cep.s(q)

    def reset(self, q: int):
        self.cep.reset(q)

    def measure(self, q: int, c: int):
        return self.cep.measure(q, c)

    def propagate_faults(self, icomb: Tuple[int], error: Tuple[str]):
        """Insert a set of faults and propagate through the loaded circuit (one kernel call)."""
        if self.encoded_circ is None:
            raise Exception("no circuit loaded")
        return self.cep.propagate(list(icomb), list(error))

    def propagate_many(self, icombs: Sequence[Sequence[int]], errors: Sequence[Sequence[str]]):
        """Batched form of ``propagate_faults``: one kernel launch for all fault lists."""
        if self.encoded_circ is None:
            raise Exception("no circuit loaded")
        return self.cep.propagate_many([list(i) for i in icombs], [list(e) for e in errors])

    def get_qubit_array(self):
        return self.cep.get_qubit_array()

    def get_bit_array(self):
        return self.cep.get_cbits()

    def get_error(self):
        return self.cep.get_qubits()


class ErrorPropagator:
    """Factory with the reference's ``eptype`` switch.  "auto" and "c" give the compiled
    propagator; "py" (the reference's pure-Python twin) does not exist here by design."""

    def __new__(cls, eptype: str = "auto", qreg_size: int = 1, creg_size: int = 1):
        if eptype in ("auto", "c"):
            if C_ERROR_PROPAGATOR:
                return CErrorPropagator(qreg_size, creg_size)
            raise ImportError("C/C++ ErrorPropagator extension not loaded.")
        if eptype == "py":
            raise NotImplementedError("no pure-Python propagator in the B200 build (no CPU fallback)")
        raise NotImplementedError("Unsupported error propagator type")
