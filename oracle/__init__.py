"""TEST INFRASTRUCTURE ONLY.  Nothing under ``qiskit_qec_b200/`` may import this package;
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs do."""
