"""Host-side mirror of qiskit-qec's ``analysis`` drivers for the accelerated path
(reference ``src/qiskit_qec/analysis/{extensions,errorpropagator,cerrorpropagator,
faultenumerator,faultsampler}.py``)."""

from .errorpropagator import CErrorPropagator, ErrorPropagator  # noqa: F401
from .faultenumerator import FaultEnumerator  # noqa: F401
from .faultsampler import FaultSampler  # noqa: F401
