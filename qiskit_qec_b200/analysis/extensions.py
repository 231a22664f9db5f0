"""Analysis extensions: feature flags set by try-import, as the reference does
(reference ``src/qiskit_qec/analysis/extensions.py:20-96``).  The extension here is the
B200-native ``_c_analysis`` (pybind11 over ``libqecb200.so``); there is no Python fallback
behind a False flag -- the drivers raise instead."""

import logging

logger = logging.getLogger(__name__)

try:
    from qiskit_qec_b200.analysis._c_analysis import (  # noqa: F401
        _CErrorPropagator, _CFaultEnumerator, _CFaultSampler, _c_isotropic, _c_minimum_distance,
        _c_minimum_distance_by_tests, _c_rank)

    C_ERROR_PROPAGATOR = C_FAULT_ENUMERATOR = C_FAULT_SAMPLER = True
    C_MIN_DISTANCE = C_MIN_DISTANCE_BY_TESTS = C_RANK = C_ISOTROPIC = True
except ImportError as import_error:  # pragma: no cover - build problem
    logger.exception("from qiskit_qec_b200.analysis._c_analysis import ... failed, raising %s",
                     import_error)
    C_ERROR_PROPAGATOR = C_FAULT_ENUMERATOR = C_FAULT_SAMPLER = False
    C_MIN_DISTANCE = C_MIN_DISTANCE_BY_TESTS = C_RANK = C_ISOTROPIC = False
