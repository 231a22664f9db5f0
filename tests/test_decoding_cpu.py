"""CPU tests of the decoding-graph host step (qiskit_qec_b200/decoding.py):
``string2nodes`` against the reference's own vectors (test/code_circuits/test_surface_codes.py:
28-147, test_rep_codes.py:124-160), and the reference's single-error property
(test_rep_codes.py:70-120, 210-250: no single fault creates exactly one node, none spans more
than two rounds) on the order-1 streams of the compiled reference enumerator."""

import pytest

from helpers import noise_inputs
from qiskit_qec_b200.circuits import repetition_code_circuit, surface_code_circuit
from qiskit_qec_b200.decoding import (DecodingGraphNode as N, RepetitionCodeNodes, SurfaceCodeNodes,
                                      graph_from_fault_paths, outcome_to_string)
from qiskit_qec_b200.noise import standard_depolarizing_model

STRINGS = ["000000000 0000 0000", "000010000 0000 0110", "000010000 0110 0000",
           "000000001 0000 0000", "100000000 0000 0000"]
WANT = {
    "x": [[],
          [N(time=1, qubits=[0, 1, 3, 4], index=1), N(time=1, qubits=[4, 5, 7, 8], index=2)],
          [N(time=0, qubits=[0, 1, 3, 4], index=1), N(time=0, qubits=[4, 5, 7, 8], index=2)],
          [N(is_logical=True, is_boundary=True, qubits=[0, 3, 6], index=0), N(time=1, qubits=[0, 1, 3, 4], index=1)],
          [N(is_logical=True, is_boundary=True, qubits=[2, 5, 8], index=1), N(time=1, qubits=[4, 5, 7, 8], index=2)]],
    "z": [[],
          [N(time=0, qubits=[1, 4, 2, 5], index=1), N(time=0, qubits=[3, 6, 4, 7], index=2)],
          [N(time=1, qubits=[1, 4, 2, 5], index=1), N(time=1, qubits=[3, 6, 4, 7], index=2)],
          [N(is_logical=True, is_boundary=True, qubits=[0, 1, 2], index=0), N(time=1, qubits=[0, 3], index=0)],
          [N(is_logical=True, is_boundary=True, qubits=[8, 7, 6], index=1), N(time=1, qubits=[5, 8], index=3)]],
}


@pytest.mark.parametrize("basis", ["x", "z"])
def test_surface_string2nodes_reference_vectors(basis):
    code = SurfaceCodeNodes(3, 1, basis=basis)
    for string, want in zip(STRINGS, WANT[basis]):
        assert code.string2nodes(string) == want, string


def test_repetition_string2nodes_reference_vectors():
    code = RepetitionCodeNodes(5, 0, resets=True)
    assert code.string2nodes("00001") == [N(is_logical=True, is_boundary=True, qubits=[0], index=0),
                                          N(time=0, qubits=[0, 1], index=0)]
    code = RepetitionCodeNodes(3, 2)
    assert code.string2nodes("0 0  01 00 01", logical="0") == code.string2nodes("1 1  01 00 01", logical="1")


def test_node_equality_rule():
    assert N(time=1, qubits=[1, 0], index=0) == N(time=1, qubits=[0, 1], index=0)     # qubits as a set
    assert N(time=1, qubits=[0, 1], index=0) != N(time=2, qubits=[0, 1], index=0)
    assert N(is_logical=True, qubits=[0], index=0, time=3) == N(is_logical=True, qubits=[0], index=0)
    with pytest.raises(ValueError):
        N(index=0, qubits=[0])


CASES = [("surf3z", lambda: surface_code_circuit(3, 2), lambda: SurfaceCodeNodes(3, 2)),
         ("surf3x", lambda: surface_code_circuit(3, 1, basis="x"), lambda: SurfaceCodeNodes(3, 1, basis="x")),
         ("surf5z", lambda: surface_code_circuit(5, 3), lambda: SurfaceCodeNodes(5, 3)),
         ("rep3", lambda: repetition_code_circuit(3, 2), lambda: RepetitionCodeNodes(3, 2)),
         ("rep5r", lambda: repetition_code_circuit(5, 3, resets=True), lambda: RepetitionCodeNodes(5, 3, resets=True))]


@pytest.mark.parametrize("name,make_circ,make_code", CASES)
def test_single_faults_create_node_pairs(ref, name, make_circ, make_code):
    circ, code = make_circ(), make_code()
    assert list(circ.creg_sizes) == code.reg_sizes
    idx, labels, paulis = noise_inputs(circ, standard_depolarizing_model())
    fe = ref._CFaultEnumerator(1, circ.nq, circ.nc, circ.ops, idx, labels, paulis)
    paths = []
    while not fe.done():
        paths += fe.enumerate(10000)
    for _, _, _, outcome in paths:
        nodes = code.string2nodes(outcome_to_string(outcome, code.reg_sizes))
        assert len(nodes) in (0, 2)
        ts = [n.time for n in nodes if not n.is_logical]
        assert not ts or max(ts) - min(ts) <= 1
    g = graph_from_fault_paths(code, paths)
    # every hyperedge of an order-1 graph is one edge; every edge joins two distinct known nodes
    assert all(len(h) == 1 for h in g.hyperedges) and len(g.hyperedges) == len(g.edges)
    assert all(0 <= a < b < len(g.nodes) for a, b in g.edges)
