"""CPU tests of the pieces around the hot path that need no GPU:

  * ``_c_rank`` / ``_c_isotropic`` (reference ``analysis/bindings/analysis_bindings.cpp:56-61``,
    host linear algebra as in the reference) against the compiled reference (``oracle/_ref``) on
    the reference's own code tables and on random inputs; the two distance functions of the same
    block run on the GPU and are tested in ``tests/test_gpu_distance.py``;
  * ``lower()`` fed by a qiskit-shaped circuit -- a stub ``qiskit.converters``
    with duck-typed DAG nodes (``.name .qargs .cargs .op._label``), since qiskit itself is not
    installed here: op order, qubit / clbit flattening, labels, the unknown-op exception and the
    dropped conditions of reference ``analysis/cerrorpropagator.py:51-83``.
"""

import sys
import types

import numpy as np
import pytest


def _strarray(label):
    n = len(label)
    vec = [0] * (2 * n)
    for i, p in enumerate(label):
        if p in "xy":
            vec[i] = 1
        if p in "zy":
            vec[i + n] = 1
    return vec


CODES = {  # reference test/distance/test_minimum_distance_extension.py: stabilizers, distance
    "steane": (["iiixxxx", "ixxiixx", "xixixix", "iiizzzz", "izziizz", "ziziziz"], 3),
    "four": (["xxxx", "zzzz"], 2),
    "five": (["xzzxi", "ixzzx", "xixzz", "zxixz"], 3),
    "ten": (["xzizixizzi", "iyizzyiizz", "izxiiyiyzx", "izzyiyziiz", "iizzyyzxzx", "izzizzxxzz",
             "iziiizzzyy", "zzzzzziiii"], 4),
    "ten_over": (["xzizixizzi", "iyizzyiizz", "izxiiyiyzx", "izzyiyziiz", "iizzyyzxzx",
                  "izzizzxxzz", "iziiizzzyy", "zzzzzziiii", "zizzzizzyy", "ziiziixxzz"], 4),
}


@pytest.fixture(scope="module")
def ours():
    from qiskit_qec_b200.analysis import _c_analysis
    return _c_analysis


@pytest.mark.parametrize("name", sorted(CODES))
def test_rank_and_isotropic_on_reference_tables(ours, ref, name):
    paulis, _ = CODES[name]
    stab = [_strarray(p) for p in paulis]
    assert ours._c_isotropic(stab) is True and ref._c_isotropic(stab) is True
    assert ours._c_rank(stab) == ref._c_rank(stab)


def test_rank_isotropic_random_differential(ours, ref):
    rng = np.random.default_rng(2)
    for trial in range(200):
        rows, cols = int(rng.integers(1, 9)), 2 * int(rng.integers(1, 40))
        m = rng.integers(0, 2, size=(rows, cols)).astype(int).tolist()
        assert ours._c_rank(m) == ref._c_rank(m), trial
        assert ours._c_isotropic(m) == ref._c_isotropic(m), trial


def test_minimum_distance_needs_the_gpu(ours):
    """The brute-force distance functions run on the GPU (csrc/qec_distance.cu; parity tests in
    tests/test_gpu_distance.py): without a device they fail loudly, there is no host version."""
    from qiskit_qec_b200 import _capi
    if _capi.device_count() > 0:
        pytest.skip("a CUDA device is present")
    stab = [_strarray(p) for p in CODES["four"][0]]
    with pytest.raises(RuntimeError, match="no CUDA device"):
        ours._c_minimum_distance(stab, stab)
    with pytest.raises(RuntimeError, match="no CUDA device"):
        ours._c_minimum_distance_by_tests(stab, [_strarray("xxii")], [_strarray("zzii")])


# ------------------------------------------------------------------------- 8(f1): lower()
class _Bit:
    def __init__(self, name):
        self.name = name

    def __repr__(self):
        return self.name


class _Op:
    def __init__(self, name, label=None, condition=None):
        self.name, self._label, self.condition = name, label, condition


class _Node:
    def __init__(self, name, qargs=(), cargs=(), label=None, condition=None):
        self.name, self.qargs, self.cargs = name, tuple(qargs), tuple(cargs)
        self.op = _Op(name, label, condition)


class _Circuit:
    """What lower() reads of a QuantumCircuit: .qubits, .clbits, .cregs, .name, and (through the
    stub converter) the op nodes in topological order."""

    def __init__(self, n_q, creg_sizes, name="stub"):
        self.qubits = [_Bit(f"q{i}") for i in range(n_q)]
        self.cregs, self.clbits = [], []
        for r, size in enumerate(creg_sizes):
            reg = [_Bit(f"c{r}_{i}") for i in range(size)]
            self.cregs.append(reg)
            self.clbits += reg
        self.name = name
        self.nodes = []

    def add(self, name, q=(), c=(), label=None, condition=None):
        self.nodes.append(_Node(name, [self.qubits[i] for i in q], [self.clbits[i] for i in c],
                                label, condition))


class _Dag:
    def __init__(self, circ):
        self._nodes = list(circ.nodes)

    def topological_op_nodes(self):
        return iter(self._nodes)


@pytest.fixture
def stub_qiskit(monkeypatch):
    qiskit = types.ModuleType("qiskit")
    conv = types.ModuleType("qiskit.converters")
    conv.circuit_to_dag = _Dag
    qiskit.converters = conv
    monkeypatch.setitem(sys.modules, "qiskit", qiskit)
    monkeypatch.setitem(sys.modules, "qiskit.converters", conv)
    return qiskit


def _reference_test_circuit():
    """the 7-op, 3-qubit circuit of reference test/fault_enumerator/test_faultenumerator.py:14-30:
    two classical registers, a labelled identity"""
    qc = _Circuit(3, [2, 1])
    qc.add("h", [0])
    qc.add("cx", [0, 1])
    qc.add("cx", [1, 2])
    qc.add("id", [0], label="special")
    qc.add("measure", [0], [0])
    qc.add("measure", [1], [1])
    qc.add("measure", [2], [2])
    return qc


def test_lower_follows_cerrorpropagator_load_circuit(stub_qiskit):
    from qiskit_qec_b200.analysis.lowering import lower
    qc = _reference_test_circuit()
    qc.add("barrier", [0, 1, 2])
    qc.add("x", [2], condition=("c", 1))    # c_if: the condition is dropped, the gate kept (:72)
    qc.add("reset", [1])
    enc = lower(qc)
    assert (enc.nq, enc.nc) == (3, 3) and enc.creg_sizes == [2, 1]
    assert enc.ops == [[0, 0], [5, 0, 1], [5, 1, 2], [6, 0], [8, 0, 0], [8, 1, 1], [8, 2, 2],
                       [9], [2, 2], [7, 1]]
    assert enc.labels == ["h", "cx", "cx", "special", "measure", "measure", "measure", "barrier",
                          "x", "reset"]
    bad = _reference_test_circuit()
    bad.add("sdg", [0])
    with pytest.raises(Exception, match='op "sdg" not recognized'):
        lower(bad)


def test_location_tagging_from_a_lowered_quantum_circuit(stub_qiskit):
    """faulty_ops_indices / labels / Pauli lists as reference faultenumerator.py:110-116,128-140
    builds them, including the per-label channel of IGate(label="special")."""
    from qiskit_qec_b200.analysis.lowering import lower
    from qiskit_qec_b200.locations import tag_fault_locations
    from qiskit_qec_b200.noise import PauliNoiseModel
    model = PauliNoiseModel()
    model.add_operation("h", {"x": 1, "y": 1, "z": 1})
    model.add_operation("cx", {"ix": 1, "xi": 1, "xx": 1})
    model.add_operation("special", {"y": 1})
    model.add_operation("measure", {"x": 1})
    enc = lower(_reference_test_circuit())
    idx, labels = tag_fault_locations(enc.labels, model)
    assert idx == [0, 1, 2, 3, 4, 5, 6]
    assert labels == ["h", "cx", "cx", "special", "measure", "measure", "measure"]
    types_ = model.get_pauli_error_types()
    assert [types_[lab] for lab in labels][3] == ["y"]
