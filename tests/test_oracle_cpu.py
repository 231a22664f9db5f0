"""CPU suite: pins the plain-C oracle (oracle/qec_oracle.c) against
  (a) the reference's own known-answer vectors (cited per test),
  (b) the committed golden vectors produced by the unmodified reference C++
      (tests/golden/ref_vectors.json, generator tests/golden/make_golden.py),
  (c) the live reference extension (oracle/_ref) when it has been built."""

import json
import os

import numpy as np
import pytest

from helpers import noise_inputs, random_circuit, random_fault_lists
from oracle import port
from qiskit_qec_b200.circuits import repetition_code_circuit, surface_code_circuit
from qiskit_qec_b200.noise import standard_depolarizing_model

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_vectors.json")

# reference test/fault_enumerator/test_faultenumerator_extension.py:13-34
KAT_CIRC = [[7, 0], [7, 1], [7, 2], [0, 0], [5, 0, 1], [6, 2], [5, 1, 2], [6, 0],
            [8, 0, 0], [8, 1, 1], [8, 2, 2]]
KAT_IDX = [3, 4, 5, 6, 7]
KAT_LABELS = ["h", "cx", "id", "cx2", "id"]
KAT_PAULIS = [["x", "y", "x"], ["ix", "xi", "xx"], ["x", "y"],
              ["ix", "xi", "xx", "iz", "zi", "zz"], ["x"]]
# reference test_faultenumerator_extension.py:46-83 (index, labels, errors, outcome)
KAT_PATHS = [
    (0, ["h"], ["x"], [1, 1, 1]), (1, ["h"], ["y"], [1, 1, 1]), (2, ["h"], ["x"], [1, 1, 1]),
    (3, ["cx"], ["ix"], [0, 1, 1]), (4, ["cx"], ["xi"], [1, 0, 0]), (5, ["cx"], ["xx"], [1, 1, 1]),
    (6, ["id"], ["x"], [0, 0, 1]), (7, ["id"], ["y"], [0, 0, 1]),
    (8, ["cx2"], ["ix"], [0, 0, 1]), (9, ["cx2"], ["xi"], [0, 1, 0]), (10, ["cx2"], ["xx"], [0, 1, 1]),
    (11, ["cx2"], ["iz"], [0, 0, 0]), (12, ["cx2"], ["zi"], [0, 0, 0]), (13, ["cx2"], ["zz"], [0, 0, 0]),
    (14, ["id"], ["x"], [1, 0, 0]),
]


def bits(rows):
    return np.array([[int(ch) for ch in r] for r in rows], dtype=np.uint8).reshape(len(rows), -1)


def test_kat_propagate_compiled():
    """reference test/error_propagator/test_compilederrorpropagator.py:24-35"""
    circ = [[0, 0], [5, 0, 1], [5, 1, 2], [8, 0, 0], [8, 1, 1], [8, 2, 2]]
    out = port.propagate_batch(3, 4, circ, [([0], ["x"])])
    assert out.tolist() == [[1, 1, 1, 0]]


def test_kat_cerrorpropagator_circuits():
    """reference test/error_propagator/test_cerrorpropagator.py:25-84 (hand-lowered circuits)"""
    c1 = [[0, 0], [8, 0, 0]]
    got = port.propagate_batch(1, 1, c1, [([0], ["x"]), ([0], ["y"]), ([0], ["z"]), ([1], ["x"])])
    assert got.tolist() == [[1], [1], [0], [1]]
    c3 = [[0, 0], [5, 0, 1], [5, 1, 2], [6, 0], [8, 0, 0], [8, 1, 1], [8, 2, 2]]
    got = port.propagate_batch(3, 3, c3, [([0], ["x"]), ([0], ["y"]), ([0], ["z"]), ([1], ["ix"])])
    assert got.tolist() == [[1, 1, 1], [1, 1, 1], [0, 0, 0], [0, 1, 1]]


def test_kat_interactive_frame():
    """reference test_cerrorpropagator.py:11-23: y on q0, h, cx, cx, measures ->
    frame [1,1,1,1,0,0], bits [1,1,1]; restated as a circuit with the fault on an id."""
    circ = [[6, 0], [0, 0], [5, 0, 1], [5, 1, 2], [8, 0, 0], [8, 1, 1], [8, 2, 2]]
    out, frame = port.propagate_batch(3, 3, circ, [([0], ["y"])], want_frame=True)
    assert out.tolist() == [[1, 1, 1]]
    assert frame.tolist() == [[1, 1, 1, 1, 0, 0]]


def test_kat_enumerator_extension():
    """reference test/fault_enumerator/test_faultenumerator_extension.py:10-102"""
    total, out, locs, digs = port.enumerate_range(3, 3, KAT_CIRC, 1, KAT_IDX, KAT_PAULIS,
                                                  want_paths=True)
    assert total == 15
    for r, (index, labels, errors, outcome) in enumerate(KAT_PATHS):
        assert index == r
        assert [KAT_LABELS[l] for l in locs[r]] == labels
        assert [KAT_PAULIS[l][d] for l, d in zip(locs[r], digs[r])] == errors
        assert out[r].tolist() == outcome


def test_kat_faultenumerator_propagator():
    """reference test/fault_enumerator/test_faultenumerator.py:50-84: qiskit's topological
    order of that circuit is h, cx01, id(special), cx12, m0, m1, m2 (fault indices 6 = special,
    7..9 = second cx in the expected list)."""
    circ = [[0, 0], [5, 0, 1], [6, 0], [5, 1, 2], [8, 0, 0], [8, 1, 1], [8, 2, 2]]
    idx = [0, 1, 2, 3, 4, 5, 6]
    paulis = [["x", "y", "z"], ["ix", "xi", "xx"], ["y"], ["ix", "xi", "xx"], ["x"], ["x"], ["x"]]
    want = [[1, 1, 1], [1, 1, 1], [0, 0, 0], [0, 1, 1], [1, 0, 0], [1, 1, 1], [1, 0, 0],
            [0, 0, 1], [0, 1, 0], [0, 1, 1], [1, 0, 0], [0, 1, 0], [0, 0, 1]]
    total, out = port.enumerate_range(3, 3, circ, 1, idx, paulis)
    assert total == 13 and out.tolist() == want


def test_semantics_corner_cases():
    """overwrite-not-xor measure, reset wipes, fault-before-measure, duplicate op: last wins
    (reference errorpropagator.cpp:250-254, :239-243, :296-300, :265-267)."""
    circ = [[6, 0], [8, 0, 0], [7, 0], [8, 0, 0]]
    assert port.propagate_batch(1, 1, circ, [([0], ["x"])]).tolist() == [[0]]
    assert port.propagate_batch(1, 2, [[6, 0], [8, 0, 0], [8, 0, 1]], [([1], ["x"])]).tolist() == [[1, 1]]
    assert port.propagate_batch(1, 1, [[6, 0], [8, 0, 0]], [([0, 0], ["x", "i"])]).tolist() == [[0]]
    assert port.propagate_batch(1, 1, [[6, 0], [8, 0, 0]], [([0, 0], ["i", "x"])]).tolist() == [[1]]
    # s then h: x -> y -> (h) y ; z unaffected by s
    out, fr = port.propagate_batch(1, 1, [[6, 0], [1, 0], [0, 0], [8, 0, 0]], [([0], ["x"])], True)
    assert out.tolist() == [[1]] and fr.tolist() == [[1, 1]]


def test_golden_propagate():
    gold = json.load(open(GOLD))
    for case in gold["propagate"]:
        paths = [(p[0], p[1]) for p in case["paths"]]
        out, fr = port.propagate_batch(case["nq"], case["nc"], case["ops"], paths, want_frame=True)
        np.testing.assert_array_equal(out, bits(case["outcomes"]))
        np.testing.assert_array_equal(fr, bits(case["frames"]))


def test_golden_enumerate():
    gold = json.load(open(GOLD))
    for case in gold["enumerate"]:
        total, out, locs, digs = port.enumerate_range(case["nq"], case["nc"], case["ops"],
                                                      case["order"], case["loc_op"], case["paulis"],
                                                      want_paths=True)
        assert total == case["total"]
        keep = case["index"]
        np.testing.assert_array_equal(out[keep], bits(case["outcomes"]))
        for k, r in enumerate(keep):
            assert [case["paulis"][l][d] for l, d in zip(locs[r], digs[r])] == case["errors"][k]
            assert [case["labels"][l] for l in locs[r]] == case["path_labels"][k]


def test_live_reference_differential(ref):
    """oracle port vs the compiled reference on fresh random inputs and the BASELINE circuits"""
    rng = np.random.default_rng(2024)
    for nq, nc, n_ops in [(2, 2, 25), (9, 6, 120), (49, 145, 761)]:
        circ = random_circuit(rng, nq, nc, n_ops)
        paths = random_fault_lists(rng, circ, 300, 10)
        ep = ref._CErrorPropagator(nq, nc)
        ep.load_circuit(nq, nc, circ.ops)
        want = np.array([ep.propagate(i, e) for i, e in paths], dtype=np.uint8).reshape(len(paths), nc)
        np.testing.assert_array_equal(port.propagate_batch(nq, nc, circ.ops, paths), want)
    model = standard_depolarizing_model()
    for circ, order in [(repetition_code_circuit(3, 2), 2), (surface_code_circuit(3, 2), 1)]:
        idx, labels, paulis = noise_inputs(circ, model)
        fe = ref._CFaultEnumerator(order, circ.nq, circ.nc, circ.ops, idx, labels, paulis)
        blk = []
        while not fe.done():
            blk += fe.enumerate(5000)
        total, out = port.enumerate_range(circ.nq, circ.nc, circ.ops, order, idx, paulis)
        assert total == len(blk)
        np.testing.assert_array_equal(out, np.array([t[3] for t in blk], dtype=np.uint8))


def test_philox_known_answers():
    """Random123 philox4x32-10 kat_vectors"""
    assert port.philox([0, 0, 0, 0], [0, 0]) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert port.philox([0xffffffff] * 4, [0xffffffff] * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert port.philox([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0]) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def test_computed_gap_is_geometric():
    """The sampler's computed gap (oracle/sampler_port.c gap_len == qec_gap_len on the device):
    known answers, clamping, and P(gap >= k) = (1 - p)^k over a regular sweep of the 32-bit
    draw (2^22 points, so the empirical tail is exact to ~2^-22 plus the polynomial's error)."""
    import ctypes as C
    from oracle import port
    lib = port.lib()
    lib.orc_gap_len.restype = C.c_uint32
    lib.orc_gap_len.argtypes = [C.c_uint32, C.c_float]
    for p in (1e-3, 0.05, 0.5):
        c = np.float32(-1.0 / np.log2(1.0 - p))
        r = (np.arange(1 << 22, dtype=np.uint64) << 10).astype(np.uint32) | np.uint32(0x155)
        g = np.array([lib.orc_gap_len(int(x), c) for x in r[::64]], dtype=np.int64)
        # exact answer from double arithmetic; the float path may differ by one slot at a boundary
        u = ((r[::64] | 1).astype(np.float64)) / 2.0 ** 32
        want = np.floor(np.log(u) / np.log1p(-p))
        assert np.abs(g - want).max() <= 1 and (g != want).mean() < 2e-3
        ks = np.array([0, 1, 2, 5, 10, 100, 1000])
        tail = np.array([(g >= k).mean() for k in ks])
        assert np.allclose(tail, (1 - p) ** ks, atol=4 * np.sqrt(0.25 / len(g)))
    assert lib.orc_gap_len(0xFFFFFFFF, np.float32(692.8)) == 0           # u ~ 1: no gap
    assert lib.orc_gap_len(0, np.float32(6.9e8)) == 1 << 24               # clamped
    assert lib.orc_gap_len(0x30000000, np.float32(1.0)) == 2              # u = 3/16: -log2 u = 2.415
    assert lib.orc_gap_len(0x30000000, np.float32(10.0)) == 24
    assert lib.orc_gap_len(12345, np.float32(0.0)) == 0                   # p = 1: every slot
