"""CPU checks of the C-ABI library: it loads, exports every symbol include/qec_b200.h
declares, validates inputs like the reference, and refuses to compute without a GPU."""

import ctypes as C
import os
import re

import pytest

from qiskit_qec_b200 import _capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "qec_b200.h")).read()
    declared = set(re.findall(r"\b(qec_[a-z0-9_]+)\s*\(", header))
    assert declared == set(_capi.SYMBOLS)
    lib = _capi.lib()
    for name in declared:
        assert hasattr(lib, name), name
    assert b"sm_100a" in lib.qec_version()


def test_no_cpu_fallback():
    if _capi.device_count() > 0:
        pytest.skip("a GPU is visible")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _capi.Circuit(1, 1, [[0, 0], [8, 0, 0]])


@pytest.mark.parametrize("rows,nq,nc,msg", [
    ([[0, 0, 1]], 2, 1, "bad 1-qubit instruction"),
    ([[8, 0]], 1, 1, "bad measure instruction"),
    ([[5, 0]], 2, 1, "bad cx instruction"),
    ([[9, 0]], 1, 1, "bad barrier instruction"),
    ([[12, 0]], 1, 1, "bad opcode"),
    ([[0, 3]], 2, 1, "qubit index out of range"),
    ([[5, 0, 2]], 2, 1, "qubit index out of range"),
    ([[8, 0, 1]], 1, 1, "bit index out of range"),
])
def test_load_circuit_validation_messages(rows, nq, nc, msg):
    """same conditions and messages as reference errorpropagator.cpp:112-148 (logic_error ->
    RuntimeError); validation happens before any device is touched"""
    with pytest.raises(RuntimeError, match=msg):
        _capi.Circuit(nq, nc, rows)


def test_pauli_code():
    assert [_capi.pauli_code(p) for p in ["i", "x", "z", "y", "ix", "xi", "yz", "zz", ""]] == \
        [0, 1, 2, 3, 4, 1, 3 | (2 << 2), 2 | (2 << 2), 0]
