import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def ref():
    """The unmodified reference C++ extension (oracle/_ref), or skip."""
    from oracle import ref as _ref

    if not _ref.available():
        pytest.skip("oracle/_ref not built (run oracle/build_ref.sh where /root/reference exists)")
    return _ref.module()


@pytest.fixture(scope="session")
def port():
    """The CPU restatement (oracle/): the checker, never the thing under test."""
    from oracle import port as _port

    _port.lib()
    return _port


@pytest.fixture(scope="session")
def capi():
    """The ctypes binding of libqecb200.so with a usable GPU, or fail loudly."""
    from qiskit_qec_b200 import _capi

    _capi.lib()
    if _capi.device_count() < 1:
        pytest.fail("no CUDA device visible to libqecb200.so (gpu-marked test on a CPU box?)")
    return _capi
