"""GPU parity of the brute-force distance functions: ``_c_minimum_distance`` and
``_c_minimum_distance_by_tests`` (csrc/qec_distance.cu through the C ABI ``qec_minimum_distance`` /
``qec_minimum_distance_by_tests``) against the compiled reference (``oracle/_ref``: reference
``analysis/intern/distance.cpp:17-53``, ``:71-159``, ``:172-202``) on the reference's own code
tables (``test/distance/test_minimum_distance_extension.py``), on subsystem codes (gauge group
larger than the stabilizer), on k = 0 stabilizer states, and on random commuting sets."""

import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _strarray(label):
    n = len(label)
    vec = [0] * (2 * n)
    for i, p in enumerate(label):
        if p in "xy":
            vec[i] = 1
        if p in "zy":
            vec[i + n] = 1
    return vec


CODES = {  # reference test/distance/test_minimum_distance_extension.py: stabilizers, distance
    "steane": (["iiixxxx", "ixxiixx", "xixixix", "iiizzzz", "izziizz", "ziziziz"], 3),
    "four": (["xxxx", "zzzz"], 2),
    "five": (["xzzxi", "ixzzx", "xixzz", "zxixz"], 3),
    "ten": (["xzizixizzi", "iyizzyiizz", "izxiiyiyzx", "izzyiyziiz", "iizzyyzxzx", "izzizzxxzz",
             "iziiizzzyy", "zzzzzziiii"], 4),
    "ten_over": (["xzizixizzi", "iyizzyiizz", "izxiiyiyzx", "izzyiyziiz", "iizzyyzxzx",
                  "izzizzxxzz", "iziiizzzyy", "zzzzzziiii", "zizzzizzyy", "ziiziixxzz"], 4),
}


@pytest.fixture(scope="module")
def ours():
    from qiskit_qec_b200.analysis import _c_analysis
    return _c_analysis


@pytest.mark.parametrize("name", sorted(CODES))
def test_minimum_distance_reference_kats(ours, ref, name):
    paulis, d = CODES[name]
    stab = [_strarray(p) for p in paulis]
    assert ours._c_minimum_distance(stab, stab) == d == ref._c_minimum_distance(stab, stab)
    for mw in (1, 2, 3):
        assert ours._c_minimum_distance(stab, stab, mw) == ref._c_minimum_distance(stab, stab, mw)


def test_minimum_distance_subsystem_and_state(ours, ref):
    # Bacon-Shor [[9,1,4,3]]: stabilizer = pairs of rows / columns, gauge = two-body XX / ZZ
    def op(kind, qubits):
        s = ["i"] * 9
        for q in qubits:
            s[q] = kind
        return _strarray("".join(s))
    stab = [op("x", [0, 1, 2, 3, 4, 5]), op("x", [3, 4, 5, 6, 7, 8]),
            op("z", [0, 1, 3, 4, 6, 7]), op("z", [1, 2, 4, 5, 7, 8])]
    gauge = [op("x", [r * 3 + c, (r + 1) * 3 + c]) for r in range(2) for c in range(3)]
    gauge += [op("z", [r * 3 + c, r * 3 + c + 1]) for r in range(3) for c in range(2)]
    want = ref._c_minimum_distance(stab, gauge, 4)
    assert want == 3
    assert ours._c_minimum_distance(stab, gauge, 4) == want
    assert ours._c_minimum_distance(stab, gauge, 2) == ref._c_minimum_distance(stab, gauge, 2) == 0
    # k = 0 (a stabilizer state): the reference then looks for the lightest element of V
    ghz = [_strarray("xxx"), _strarray("zzi"), _strarray("izz")]
    assert ours._c_minimum_distance(ghz, ghz, 3) == ref._c_minimum_distance(ghz, ghz, 3)


def test_minimum_distance_random_differential(ours, ref):
    rng = np.random.default_rng(2)
    done = 0
    while done < 40:
        n = int(rng.integers(3, 8))
        gens = []
        for _ in range(int(rng.integers(1, n + 1))):
            for _try in range(50):
                v = rng.integers(0, 2, size=2 * n).astype(int).tolist()
                if any(v) and ours._c_isotropic(gens + [v]):
                    gens.append(v)
                    break
        if not gens:
            continue
        assert ours._c_minimum_distance(gens, gens, 4) == ref._c_minimum_distance(gens, gens, 4), gens
        # a gauge group that contains the stabilizer (extra rows need not commute)
        gauge = gens + [rng.integers(0, 2, size=2 * n).astype(int).tolist() for _ in range(int(rng.integers(0, 3)))]
        assert ours._c_minimum_distance(gens, gauge, 3) == ref._c_minimum_distance(gens, gauge, 3), (gens, gauge)
        done += 1


def test_minimum_distance_errors_match_reference(ours, ref):
    bad = [_strarray("xi"), _strarray("zi")]
    with pytest.raises(RuntimeError, match="isotropic"):
        ours._c_minimum_distance(bad, [_strarray("xi")])
    with pytest.raises(RuntimeError, match="isotropic"):
        ref._c_minimum_distance(bad, [_strarray("xi")])


def test_minimum_distance_by_tests_matches_reference(ours, ref):
    stab = [_strarray(p) for p in CODES["steane"][0]]
    xl, zl = [_strarray("xxxxxxx")], [_strarray("zzzzzzz")]
    assert ours._c_minimum_distance_by_tests(stab, xl, zl) == 3 == ref._c_minimum_distance_by_tests(stab, xl, zl)
    assert ours._c_minimum_distance_by_tests(stab, xl, zl, 2) == ref._c_minimum_distance_by_tests(stab, xl, zl, 2) == 0
    stab = [_strarray(p) for p in CODES["five"][0]]
    xl, zl = [_strarray("xxxxx")], [_strarray("zzzzz")]
    assert ours._c_minimum_distance_by_tests(stab, xl, zl) == ref._c_minimum_distance_by_tests(stab, xl, zl) == 3
    # every weight separately, several logicals, an empty list of Z logicals
    stab = [_strarray(p) for p in CODES["ten"][0]]
    rng = np.random.default_rng(5)
    logs = [rng.integers(0, 2, size=20).astype(int).tolist() for _ in range(3)]
    for mw in range(1, 6):
        assert ours._c_minimum_distance_by_tests(stab, logs, [], mw) == ref._c_minimum_distance_by_tests(stab, logs, [], mw), mw
        assert ours._c_minimum_distance_by_tests(stab, logs[:1], logs[1:], mw) == \
            ref._c_minimum_distance_by_tests(stab, logs[:1], logs[1:], mw), mw


def test_larger_code_through_the_c_abi(capi):
    """The rotated surface code d = 5 (25 qubits, 24 generators): distance 5 needs all weights up
    to 5 (C(25,5) * 3^5 = 1.3e7 errors at the last one) -- seconds of reference CPU time, one launch
    per weight here.  Called through the C ABI with flat uint8 matrices."""
    d = 5
    def q(r, c):
        return r * d + c
    def kind(r, c):
        return "x" if (r + c) % 2 == 0 else "z"
    stabs = []
    for r in range(d - 1):
        for c in range(d - 1):
            stabs.append((kind(r, c), [q(r, c), q(r, c + 1), q(r + 1, c), q(r + 1, c + 1)]))
    for c in range(d - 1):  # two-body X checks on the top / bottom rows, Z checks on the side columns
        if kind(0, c) == "z":
            stabs.append(("x", [q(0, c), q(0, c + 1)]))
        if kind(d - 2, c) == "z":
            stabs.append(("x", [q(d - 1, c), q(d - 1, c + 1)]))
    for r in range(d - 1):
        if kind(r, 0) == "x":
            stabs.append(("z", [q(r, 0), q(r + 1, 0)]))
        if kind(r, d - 2) == "x":
            stabs.append(("z", [q(r, d - 1), q(r + 1, d - 1)]))
    n = d * d
    mat = np.zeros((len(stabs), 2 * n), dtype=np.uint8)
    for i, (kind, qs) in enumerate(stabs):
        for x in qs:
            mat[i, x + (0 if kind == "x" else n)] = 1
    lib = capi.lib()
    dist = C.c_int(-1)
    p = mat.ctypes.data_as(C.c_void_p)
    rc = lib.qec_minimum_distance(C.c_int(0), C.c_int(n), C.c_int(len(stabs)), p, C.c_int(len(stabs)), p,
                                  C.c_int(6), C.byref(dist))
    assert rc == 0, lib.qec_last_error()
    # commuting checks of a distance-5 code with one logical qubit
    assert len(stabs) == n - 1
    assert dist.value == 5
    rc = lib.qec_minimum_distance(C.c_int(0), C.c_int(n), C.c_int(len(stabs)), p, C.c_int(len(stabs)), p,
                                  C.c_int(4), C.byref(dist))
    assert rc == 0 and dist.value == 0


def test_distance_limits_and_argument_errors(ours, capi):
    """What the C ABI refuses: empty input, ragged rows (binding), weights above 16, more stabilizer
    generators than the 64-bit syndromes of distance_test hold."""
    lib = capi.lib()
    d = C.c_int(0)
    z = np.zeros((1, 4), dtype=np.uint8)
    p = z.ctypes.data_as(C.c_void_p)
    assert lib.qec_minimum_distance(C.c_int(0), C.c_int(0), C.c_int(1), p, C.c_int(1), p, C.c_int(3), C.byref(d)) == -2
    with pytest.raises(ValueError):
        ours._c_minimum_distance([[1, 0, 0, 0], [0, 1, 0]], [[1, 0, 0, 0]])
    with pytest.raises(ValueError):
        ours._c_minimum_distance([], [])
    # 70 commuting generators (Z on one qubit each, 70 qubits): distance_test needs m < 63
    n = 70
    stab = np.zeros((n, 2 * n), dtype=int)
    for i in range(n):
        stab[i, n + i] = 1
    with pytest.raises(RuntimeError, match="n-k must be < 63"):
        ours._c_minimum_distance_by_tests(stab.tolist(), [[1] * n + [0] * n], [], 2)
    # minimum_distance on the same stabilizer state (k = 0): the lightest element of V has weight 1
    assert ours._c_minimum_distance(stab.tolist(), stab.tolist(), 2) == 1
