#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY -- builds the UNMODIFIED reference C++ (qiskit-qec
# analysis/intern/*.cpp + analysis/bindings/analysis_bindings.cpp) from where the
# sources lie under /root/reference into oracle/_ref/ (git-ignored, travels to the
# GPU box with the snapshot). No reference source is copied into this repo.
# Flags mirror the reference's CMake (CMakeLists.txt:34,41; analysis/CMakeLists.txt:34).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${QEC_REFERENCE_ROOT:-/root/reference}/src/qiskit_qec/analysis"
OUT="$HERE/_ref"
PY="${PYTHON:-python}"
if [ ! -d "$REF" ]; then
  echo "build_ref: reference sources not found at $REF (fine on the GPU box: prebuilt _ref travels)" >&2
  exit 0
fi
mkdir -p "$OUT"
SUFFIX="$($PY -c 'import sysconfig;print(sysconfig.get_config_var("EXT_SUFFIX"))')"
TARGET="$OUT/_c_analysis$SUFFIX"
if [ -f "$TARGET" ] && [ -z "${QEC_FORCE_REBUILD:-}" ]; then
  echo "build_ref: $TARGET up to date"; exit 0
fi
g++ -O3 -std=c++20 -shared -fPIC -fno-strict-aliasing -w \
    -I"$REF" $($PY -m pybind11 --includes) \
    "$REF"/intern/*.cpp "$REF"/bindings/analysis_bindings.cpp \
    -o "$TARGET"
echo "build_ref: built $TARGET"
