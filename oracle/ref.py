"""TEST INFRASTRUCTURE ONLY -- loader for the UNMODIFIED reference C++ built by
``oracle/build_ref.sh`` into ``oracle/_ref/`` (reference
``src/qiskit_qec/analysis/bindings/analysis_bindings.cpp:12-54``: ``_CErrorPropagator``,
``_CFaultEnumerator``, ``_CFaultSampler``).  The built module travels to the GPU box with
the repo snapshot; ``/root/reference`` itself is never read at run time."""

from __future__ import annotations

import glob
import importlib.util
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_mod = None


def available() -> bool:
    return bool(glob.glob(os.path.join(_HERE, "_ref", "_c_analysis*.so")))


def module():
    """The reference ``_c_analysis`` extension module (raises ImportError if not built)."""
    global _mod
    if _mod is None:
        hits = glob.glob(os.path.join(_HERE, "_ref", "_c_analysis*.so"))
        if not hits:
            raise ImportError("oracle/_ref/_c_analysis*.so missing: run oracle/build_ref.sh")
        spec = importlib.util.spec_from_file_location("_c_analysis", hits[0])
        _mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(_mod)
    return _mod
