"""qiskit_qec_b200 -- B200-native (sm_100a) fault-path analysis behind qiskit-qec's
``_CErrorPropagator`` / ``_CFaultEnumerator`` / ``_CFaultSampler`` extension API.

Scope: Pauli-frame propagation of faults through a stabilizer
syndrome-extraction circuit, as driven by ``FaultSampler.sample(N)`` and
``FaultEnumerator.generate()/generate_blocks()``.  The compute lives in
``libqecb200.so`` (CUDA kernels + C ABI, ``include/qec_b200.h``); this package is the
host-side mirror of the reference interface.  There is no CPU fallback.
"""

__version__ = "0.1.0"
