"""CPU checks of the host-side mirror: circuit emitters, noise model, location tagging."""

import numpy as np
import pytest

from qiskit_qec_b200.circuits import (arc_circuit, heavy_hex_27_arc_layout,
                                      repetition_code_circuit, surface_code_circuit)
from qiskit_qec_b200.locations import tag_fault_locations
from qiskit_qec_b200.noise import PauliNoiseModel, standard_depolarizing_model


def test_baseline_circuit_shapes():
    """sizes stated in SURVEY.md section 8 header"""
    c = repetition_code_circuit(3, 2)
    assert (c.nq, c.nc, len(c)) == (5, 7, 15)
    assert c.ops == [[5, 2, 0], [5, 3, 1], [5, 3, 0], [5, 4, 1], [8, 0, 0], [8, 1, 1], [5, 2, 0],
                     [5, 3, 1], [5, 3, 0], [5, 4, 1], [8, 0, 2], [8, 1, 3], [8, 2, 4], [8, 3, 5],
                     [8, 4, 6]]
    c = surface_code_circuit(5, 5)
    assert (c.nq, c.nc, len(c)) == (49, 145, 761)
    assert c.gate_counts() == {"h": 120, "cx": 400, "measure": 145, "reset": 96}
    assert c.algorithmic_bit_ops() == 1377
    assert c.creg_sizes == [12, 12] * 5 + [25]
    c = surface_code_circuit(11, 11)
    assert (c.nq, c.nc, len(c)) == (241, 1441, 8801) and c.algorithmic_bit_ops() == 16161
    links, sched, color = heavy_hex_27_arc_layout()
    c = arc_circuit(links, 3, "zx", sched, color)
    assert c.nq == 21 and c.nc == 3 * 11 + 10
    for layer in sched:  # a layer never touches a qubit twice
        used = [q for pair in layer for q in pair]
        assert len(used) == len(set(used))
    assert all(color[a] != color[b] for a, _, b in links)


def test_surface_code_x_basis_and_logical_one():
    c = surface_code_circuit(3, 1, basis="x", logical="1")
    assert [r[0] for r in c.ops[:9]] == [0] * 9 and [r[0] for r in c.ops[9:12]] == [4] * 3
    assert repetition_code_circuit(3, 1, logical="1").ops[:3] == [[2, 2], [2, 3], [2, 4]]


def test_noise_model_mirrors_reference():
    """reference test/test_paulinoisemodel.py: normalisation, scale factors, errors"""
    m = PauliNoiseModel()
    m.add_operation("cx", {"ix": 1, "xi": 1, "xx": 1})
    m.add_operation("h", {"x": 2, "z": 6})
    assert m.get_pauli_weight("cx", "ix") == pytest.approx(1 / 3)
    assert m.get_pauli_weight("h", "z") == pytest.approx(0.75)
    assert m.get_pauli_weight("h", "y") == 0
    assert m.get_operations() == ["cx", "h"]
    assert m.get_pauli_error_types() == {"cx": ["ix", "xi", "xx"], "h": ["x", "z"]}
    with pytest.raises(Exception, match="no error probability"):
        m.get_error_probability("h")
    m.set_scale_factor("h", 2)
    m.set_scaled_error_probabilities(0.01)
    assert m.get_error_probability("h") == pytest.approx(0.02)
    assert m.get_pauli_error_probability("h", "x") == pytest.approx(0.005)
    with pytest.raises(Exception):
        m.set_error_probability("h", 1.5)
    with pytest.raises(Exception):
        m.add_operation("bad", {})
    with pytest.raises(Exception):
        m.add_operation("bad", {"x": 1, "xx": 1})
    again = PauliNoiseModel(fromdict=m.to_dict())
    assert again.to_dict() == m.to_dict()
    d = standard_depolarizing_model(1e-3)
    assert d.get_pauli_weight("cx", "zz") == pytest.approx(1 / 15)
    assert len(d.get_pauli_error_types()["cx"]) == 15


def test_location_tagging():
    c = surface_code_circuit(5, 5)
    idx, labels = tag_fault_locations(c.labels, standard_depolarizing_model())
    assert idx == list(range(761))
    m = PauliNoiseModel()
    m.add_operation("measure", {"x": 1})
    idx, labels = tag_fault_locations(c.labels, m)
    assert len(idx) == 145 and set(labels) == {"measure"}


def test_counts_from_packed_matches_qiskit_format():
    """Counts adaptor (reference utils/stim_tools.py:266-277 format): last clbit first, a space
    where the register changes, multiplicities of identical outcomes."""
    from qiskit_qec_b200.analysis.faultsampler import counts_from_packed
    rng = np.random.default_rng(5)
    nc, n, regs = 7, 75, [2, 2, 3]
    outcomes = (rng.random((n, nc)) < 0.3).astype(np.uint8)
    outcomes[10] = outcomes[3]
    packed = np.zeros((nc, (n + 31) // 32), dtype=np.uint32)
    for s in range(n):
        for c in range(nc):
            if outcomes[s, c]:
                packed[c, s // 32] |= np.uint32(1 << (s % 32))
    want = {}
    for s in range(n):
        key, prev = "", None
        reg_of = [r for r, size in enumerate(regs) for _ in range(size)]
        for c in range(nc - 1, -1, -1):
            if prev is not None and reg_of[c] != prev:
                key += " "
            key += str(int(outcomes[s, c]))
            prev = reg_of[c]
        want[key] = want.get(key, 0) + 1
    assert counts_from_packed(packed, n, regs) == want
    assert sum(counts_from_packed(packed, n, [nc]).values()) == n
    with pytest.raises(ValueError):
        counts_from_packed(packed, n, [3, 3])


def test_unpack_events16_decodes_a_synthetic_stream():
    """The host decoder of the delta-coded 16-bit event form (include/qec_b200.h,
    qec_sampler_sample_events16): entries (delta << 4 | digit) and 0xFFF0 skips per word -> events
    location << 14 | shot bit << 9 | digit through the sorted -> original location map."""
    import numpy as np
    from qiskit_qec_b200 import _capi
    L = 300
    tables = np.zeros(8 + 1 + 32 + 5 * L + 50, dtype=np.uint32)
    tables[0], tables[1], tables[3] = 32, 1, L
    tables[8 + 1 + 32 + L: 8 + 1 + 32 + 2 * L] = np.arange(L)[::-1]
    rng = np.random.default_rng(0)
    wc, ent, want, wantc = [], [], [], []
    for w in range(80):
        slots = np.sort(rng.choice(L * 32, size=int(rng.integers(0, 6)), replace=False))
        prev, n = -1, 0
        for sl in slots:
            d = int(sl) - prev - 1
            while d >= 4095:
                ent.append(0xFFF0); d -= 4095; n += 1
            dig = int(rng.integers(0, 15))
            ent.append((d << 4) | dig); n += 1; prev = int(sl)
            want.append(((L - 1 - (int(sl) >> 5)) << 14) | ((int(sl) & 31) << 9) | dig)
        wc.append(n); wantc.append(len(slots))
    assert 0xFFF0 in ent  # (skips are exercised)
    c, e = _capi.unpack_events16(np.array(wc, dtype=np.uint16), np.array(ent, dtype=np.uint16), tables)
    assert np.array_equal(c, np.array(wantc)) and np.array_equal(e, np.array(want, dtype=np.uint32))
    c, e = _capi.unpack_events16(np.zeros(3, dtype=np.uint16), np.zeros(0, dtype=np.uint16), tables)
    assert len(e) == 0 and np.array_equal(c, [0, 0, 0])


def test_rows_to_csr_and_the_csr_form_of_a_circuit():
    """`rows_to_csr` lays the reference's list of rows out as qec_circuit_create reads it; a
    Circuit also takes that pair directly and refuses offsets that do not describe the data."""
    from qiskit_qec_b200 import _capi
    rows = [[5, 0, 1], [0, 2], [9], [8, 1, 0]]
    off, data = _capi.rows_to_csr(rows)
    assert off.dtype == np.int32 and data.dtype == np.int32
    assert off.tolist() == [0, 3, 5, 6, 9]
    assert data.tolist() == [5, 0, 1, 0, 2, 9, 8, 1, 0]
    off0, data0 = _capi.rows_to_csr([])
    assert off0.tolist() == [0] and len(data0) == 0
    circ = surface_code_circuit(3, 2)
    off, data = _capi.rows_to_csr(circ.ops)
    assert len(off) == len(circ.ops) + 1 and int(off[-1]) == sum(len(r) for r in circ.ops)
    assert [data[off[i]:off[i + 1]].tolist() for i in range(len(circ.ops))] == [list(r) for r in circ.ops]
    for bad in ((np.array([1, 3], np.int32), np.zeros(3, np.int32)),        # does not start at 0
                (np.array([0, 2], np.int32), np.zeros(3, np.int32)),        # does not cover the data
                (np.array([0, 3, 2, 3], np.int32), np.zeros(3, np.int32))):  # decreasing
        with pytest.raises(ValueError):
            _capi.Circuit(2, 2, bad)


def _pack_columns(m):
    """bool [nc][F] -> uint32 [nc][ceil(F / 32)], fault f in bit f % 32 of word f // 32"""
    nc, F = m.shape
    fw = (F + 31) // 32
    pad = np.zeros((nc, fw * 32), dtype=np.uint32)
    pad[:, :F] = m
    return (pad.reshape(nc, fw, 32) << np.arange(32, dtype=np.uint32)).sum(axis=2, dtype=np.uint64).astype(np.uint32)


@pytest.mark.parametrize("seed,nc,F,density", [(1, 1, 1, 0.5), (2, 7, 127, 0.3), (3, 145, 6601, 0.04),
                                               (4, 200, 1000, 0.5), (5, 33, 32, 0.1), (6, 64, 4097, 0.01)])
def test_planner_signature_lists_match_numpy(seed, nc, F, density):
    """The host step of the linear-response planners (time-difference basis on the bit-sliced
    rows, list extraction in one or two passes) against a direct numpy statement of it."""
    from qiskit_qec_b200 import _capi
    rng = np.random.default_rng(seed)
    m = rng.random((nc, F)) < density
    # chains of "the same qubit measured again": every clbit has at most one successor
    pred = np.full(nc, -1, dtype=np.int32)
    step = max(1, nc // 6)
    for c in range(step, nc):
        if rng.random() < 0.8:
            pred[c] = c - step
    bits = _pack_columns(m)
    # stray bits past the last fault must not reach any list
    if F % 32:
        bits[:, -1] |= np.uint32((0xFFFFFFFF << (F % 32)) & 0xFFFFFFFF)
    d = m.copy()
    has = pred >= 0
    d[has] ^= m[pred[has]]
    for diff, want in ((False, m), (True, d)):
        for short in (False, True):
            off, ent, raw, dif = _capi.planner_signature_lists(bits, F, pred, diff, short)
            assert raw == int(m.sum()) and dif == int(d.sum())
            assert off[0] == 0 and off[-1] == int(want.sum()) == len(ent)
            cols = [np.nonzero(want[:, f])[0] for f in range(F)]
            np.testing.assert_array_equal(np.diff(off), [len(c) for c in cols])
            np.testing.assert_array_equal(ent, np.concatenate(cols) if len(ent) else np.zeros(0, np.uint16))


def test_planner_time_difference_basis_shortens_repeated_stabiliser_flips():
    """A fault that flips a stabiliser flips every later measurement of it: raw lists grow with
    the rounds, the time-difference lists stay at one entry."""
    from qiskit_qec_b200 import _capi
    rounds, stabs = 5, 4
    nc = rounds * stabs
    pred = np.array([c - stabs if c >= stabs else -1 for c in range(nc)], dtype=np.int32)
    m = np.zeros((nc, rounds * stabs), dtype=bool)       # fault f = (round r, stabiliser s)
    for r in range(rounds):
        for s_ in range(stabs):
            m[[q * stabs + s_ for q in range(r, rounds)], r * stabs + s_] = True
    off, ent, raw, dif = _capi.planner_signature_lists(_pack_columns(m), m.shape[1], pred, True, True)
    assert raw == stabs * sum(range(1, rounds + 1)) and dif == rounds * stabs
    np.testing.assert_array_equal(np.diff(off), np.ones(rounds * stabs, dtype=np.int64))
    np.testing.assert_array_equal(ent, np.arange(nc))
