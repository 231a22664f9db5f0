"""Fault-location tagging: which rows of an encoded circuit are faulty under a noise model.

Follows the reference drivers' ``_process_circuit`` (reference
``src/qiskit_qec/analysis/faultenumerator.py:128-140``, ``faultsampler.py:94-106``):
row ``i`` is a fault location iff its label (``node_name_label``: gate label if set, else
gate name) is one of ``model.get_operations()``.
"""

from __future__ import annotations

from typing import List, Sequence, Tuple


def tag_fault_locations(labels: Sequence[str], model) -> Tuple[List[int], List[str]]:
    """Return (faulty_ops_indices, faulty_ops_labels) in circuit order."""
    location_types = set(model.get_operations())
    idx = [i for i, lab in enumerate(labels) if lab in location_types]
    return idx, [labels[i] for i in idx]
