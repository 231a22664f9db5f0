"""SURVEY 8(f3) on the GPU: the syndrome graph built from ONE order-1 enumerator launch
(``decoding_graph_from_enumerator``) equals the graph the reference's per-fault loop builds from
the compiled reference's ``_CFaultEnumerator`` stream through the same host code: same nodes in
the same order, same edges with the same qubit lists, same hyperedges."""

import pytest

from qiskit_qec_b200.circuits import repetition_code_circuit, surface_code_circuit
from qiskit_qec_b200.decoding import (RepetitionCodeNodes, SurfaceCodeNodes, decoding_graph_from_enumerator,
                                      graph_from_fault_paths)
from qiskit_qec_b200.noise import standard_depolarizing_model

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["surf3z", "surf3x", "surf5z", "rep3", "rep5r"])
def test_graph_from_gpu_enumeration_equals_reference_loop(capi, ref, name):
    from qiskit_qec_b200.analysis import FaultEnumerator
    circ, code = {
        "surf3z": (surface_code_circuit(3, 2), SurfaceCodeNodes(3, 2)),
        "surf3x": (surface_code_circuit(3, 1, basis="x"), SurfaceCodeNodes(3, 1, basis="x")),
        "surf5z": (surface_code_circuit(5, 5), SurfaceCodeNodes(5, 5)),
        "rep3": (repetition_code_circuit(3, 2), RepetitionCodeNodes(3, 2)),
        "rep5r": (repetition_code_circuit(5, 3, resets=True), RepetitionCodeNodes(5, 3, resets=True)),
    }[name]
    fe = FaultEnumerator(circ, order=1, model=standard_depolarizing_model())
    got = decoding_graph_from_enumerator(code, fe)
    rfe = ref._CFaultEnumerator(1, circ.nq, circ.nc, circ.ops, fe.faulty_ops_indices, fe.faulty_ops_labels,
                                fe.faulty_ops_pauli_errors)
    paths = []
    while not rfe.done():
        paths += rfe.enumerate(10000)
    assert len(paths) == fe.num_paths()
    want = graph_from_fault_paths(code, paths)
    assert got.nodes == want.nodes and len(got.nodes) > 4
    assert got.edges == want.edges
    assert got.hyperedges == want.hyperedges
    # and through the driver's own generate(): " "-separated outcomes, same graph
    assert graph_from_fault_paths(code, fe.generate()).edges == want.edges
