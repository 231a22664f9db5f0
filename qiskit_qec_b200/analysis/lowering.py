"""Circuit lowering: ``QuantumCircuit`` or ``EncodedCircuit`` -> ``EncodedCircuit``.

For a qiskit ``QuantumCircuit`` this follows the reference's
``CErrorPropagator.load_circuit`` (reference ``analysis/cerrorpropagator.py:51-83``):
``circuit_to_dag`` -> ``dag.topological_op_nodes()``; op index = position in that order;
qubit / clbit index = position in ``circ.qubits`` / ``circ.clbits``; unknown op names raise
``Exception('op "..." not recognized')``; conditions are dropped.  Labels follow
``node_name_label`` (reference ``utils/dag.py:18-27``).  qiskit is imported lazily and only
for that input type.
"""

from __future__ import annotations

from qiskit_qec_b200.circuits import NAME_TO_CODE, OP_NAMES, EncodedCircuit


def node_name_label(node) -> str:
    """Gate label if set, else gate name (reference ``utils/dag.py:18-27``)."""
    label = getattr(node.op, "_label", None) or getattr(node.op, "label", None)
    return label if label is not None else node.name


def lower(circ) -> EncodedCircuit:
    if isinstance(circ, EncodedCircuit):
        return circ
    from qiskit.converters import circuit_to_dag  # only needed for QuantumCircuit inputs

    dag = circuit_to_dag(circ)
    qidx = {bit: i for i, bit in enumerate(circ.qubits)}
    cidx = {bit: i for i, bit in enumerate(circ.clbits)}
    out = EncodedCircuit(nq=len(circ.qubits), nc=len(circ.clbits), name=getattr(circ, "name", ""))
    out.creg_sizes = [len(reg) for reg in circ.cregs]
    for node in dag.topological_op_nodes():
        name = node.name
        if name not in OP_NAMES:
            raise Exception(f'op "{name}" not recognized')
        q = [qidx[b] for b in node.qargs]
        c = [cidx[b] for b in node.cargs]
        if name == "barrier":
            row = [NAME_TO_CODE[name]]
        elif name == "cx":
            row = [NAME_TO_CODE[name], q[0], q[1]]
        elif name == "measure":
            row = [NAME_TO_CODE[name], q[0], c[0]]
        else:
            row = [NAME_TO_CODE[name], q[0]]
        out.ops.append(row)
        out.labels.append(node_name_label(node))
    return out
