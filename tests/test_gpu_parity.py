"""GPU parity tests proper: CUDA path (through the C ABI) vs the unmodified reference C++
(oracle/_ref) on the same seeded inputs.  Bit-exact for everything deterministic; the
sampler is checked by replaying its exported fault lists through the reference."""

import numpy as np
import pytest

from helpers import (fault_lists_to_csr, make_enumerator, make_sampler, noise_inputs,
                     random_circuit, random_fault_lists, ref_outcome_matrix, sampler_inputs)
from qiskit_qec_b200.circuits import (arc_circuit, heavy_hex_27_arc_layout,
                                      repetition_code_circuit, surface_code_circuit)
from qiskit_qec_b200.noise import PauliNoiseModel, standard_depolarizing_model

pytestmark = pytest.mark.gpu


def _ref_propagate_all(ref, circ, paths):
    ep = ref._CErrorPropagator(circ.nq, circ.nc)
    ep.load_circuit(circ.nq, circ.nc, circ.ops)
    outs, frames = [], []
    for icomb, errs in paths:
        outs.append(ep.propagate(icomb, errs))
        frames.append(ep.get_qubit_array())
    return (np.array(outs, dtype=np.uint8).reshape(len(paths), circ.nc),
            np.array(frames, dtype=np.uint8).reshape(len(paths), 2 * circ.nq))


@pytest.mark.parametrize("seed,nq,nc,n_ops,n_paths,kmax", [
    (1, 1, 1, 5, 7, 2), (2, 3, 4, 40, 100, 4), (3, 17, 9, 300, 1000, 12),
    (4, 64, 70, 2000, 333, 12), (5, 5, 3, 0, 40, 0), (6, 130, 33, 500, 2049, 6),
])
def test_k1_random_circuits(capi, ref, seed, nq, nc, n_ops, n_paths, kmax):
    rng = np.random.default_rng(seed)
    circ = random_circuit(rng, nq, nc, n_ops)
    paths = random_fault_lists(rng, circ, n_paths, kmax if len(circ.ops) else 0)
    want, want_fr = _ref_propagate_all(ref, circ, paths)
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    bits, frame = g.propagate_batch(*fault_lists_to_csr(paths), want_frame=True)
    np.testing.assert_array_equal(capi.unpack_bits(bits, n_paths), want)
    np.testing.assert_array_equal(capi.unpack_bits(frame, n_paths), want_fr)


def test_k1_surface_d5(capi, ref):
    rng = np.random.default_rng(11)
    circ = surface_code_circuit(5, 5)
    paths = random_fault_lists(rng, circ, 4000, 12)
    want, _ = _ref_propagate_all(ref, circ, paths)
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    bits = g.propagate_batch(*fault_lists_to_csr(paths))
    np.testing.assert_array_equal(capi.unpack_bits(bits, len(paths)), want)
    # the same circuit handed over as the C ABI's own host buffers (row offsets + row data)
    g2 = capi.Circuit(circ.nq, circ.nc, capi.rows_to_csr(circ.ops))
    assert g2.n_ops == len(circ.ops)
    np.testing.assert_array_equal(g2.propagate_batch(*fault_lists_to_csr(paths)), bits)


def _ref_enum_all(ref, circ, order, idx, labels, paulis, blocksize=10000, max_paths=None):
    fe = ref._CFaultEnumerator(order, circ.nq, circ.nc, circ.ops, idx, labels, paulis)
    out = []
    while not fe.done():
        out += fe.enumerate(blocksize)
        if max_paths is not None and len(out) >= max_paths:
            break
    return out


def _check_enum_range(capi, en, circ, lo, hi, ref_paths, idx, paulis):
    bits, tally = en.range(lo, hi, want_bits=True, want_tally=True)
    got = capi.unpack_bits(bits, hi - lo)
    want = ref_outcome_matrix(ref_paths[lo:hi]) if hi > lo else np.zeros((0, circ.nc), np.uint8)
    np.testing.assert_array_equal(got, want)
    np.testing.assert_array_equal(tally[:circ.nc], want.sum(axis=0).astype(np.uint64))
    assert tally[circ.nc] == np.count_nonzero(want.any(axis=1))
    assert tally[circ.nc + 1] == hi - lo
    # host unranking agrees with the reference's (index, labels, errors)
    for r in {lo, (lo + hi) // 2, hi - 1} if hi > lo else set():
        locs, digs = en.unrank(r)
        assert ref_paths[r][0] == r
        assert ref_paths[r][2] == [paulis[l][d] for l, d in zip(locs, digs)]


@pytest.mark.parametrize("name,order", [("rep", 1), ("rep", 2), ("rep", 3), ("surf3", 1),
                                        ("surf3", 2), ("arc", 1), ("randlabels", 2)])
def test_k2_full_streams(capi, ref, name, order):
    if name == "rep":
        circ = repetition_code_circuit(3, 2)
        model = standard_depolarizing_model()
    elif name == "surf3":
        circ = surface_code_circuit(3, 2)
        model = standard_depolarizing_model()
    elif name == "arc":
        links, sched, color = heavy_hex_27_arc_layout()
        circ = arc_circuit(links, 2, "zx", sched, color)
        model = standard_depolarizing_model()
    else:
        rng = np.random.default_rng(5)
        circ = random_circuit(rng, 6, 5, 40)
        model = PauliNoiseModel()
        model.add_operation("h", {"x": 1, "z": 2})
        model.add_operation("cx", {"xi": 1, "zz": 1, "iy": 1, "yx": 3})
        model.add_operation("measure", {"x": 1})
        model.add_operation("s", {"y": 1, "x": 1, "z": 1, "i": 1})
        model.add_operation("reset", {"x": 1, "z": 1})
        model.add_operation("barrier", {"x": 1})
    idx, labels, paulis = noise_inputs(circ, model)
    ref_paths = _ref_enum_all(ref, circ, order, idx, labels, paulis)
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    en = make_enumerator(capi, g, order, idx, paulis)
    n = en.num_paths()
    assert n == len(ref_paths)
    _check_enum_range(capi, en, circ, 0, n, ref_paths, idx, paulis)
    # ragged windows: unaligned starts, single path, empty
    rng = np.random.default_rng(99)
    for _ in range(6):
        lo = int(rng.integers(0, n))
        hi = int(min(n, lo + rng.integers(0, 200)))
        _check_enum_range(capi, en, circ, lo, hi, ref_paths, idx, paulis)


def test_k2_surface_d5_order2_windows(capi, ref):
    """BASELINE cfg3 circuit: first window vs the reference stream, later windows vs the
    reference propagate() fed with host-unranked fault lists."""
    circ = surface_code_circuit(5, 5)
    model = standard_depolarizing_model()
    idx, labels, paulis = noise_inputs(circ, model)
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    en = make_enumerator(capi, g, 2, idx, paulis)
    n = en.num_paths()
    assert n == 21740940
    head = _ref_enum_all(ref, circ, 2, idx, labels, paulis, blocksize=20000, max_paths=20000)
    _check_enum_range(capi, en, circ, 0, 20000, head, idx, paulis)
    ep = ref._CErrorPropagator(circ.nq, circ.nc)
    ep.load_circuit(circ.nq, circ.nc, circ.ops)
    rng = np.random.default_rng(3)
    for lo in [n - 3000] + [int(x) for x in rng.integers(0, n - 3000, size=4)]:
        hi = lo + 3000 - int(rng.integers(0, 40))
        bits, _ = en.range(lo, hi)
        got = capi.unpack_bits(bits, hi - lo)
        for r in range(lo, hi, 7):
            locs, digs = en.unrank(r)
            want = ep.propagate([idx[l] for l in locs], [paulis[l][d] for l, d in zip(locs, digs)])
            assert got[r - lo].tolist() == want, f"path {r}"


def _decode_faults(cnt, ev, n_shots):
    """export -> per shot list of (location, digit) in event order"""
    per_shot = [[] for _ in range(n_shots)]
    for w in range(len(cnt)):
        for e in ev[w, :cnt[w]]:
            loc, bit, dig = int(e) >> 14, (int(e) >> 9) & 31, int(e) & 511
            per_shot[32 * w + bit].append((loc, dig))
    return per_shot


@pytest.mark.parametrize("name,p,n_shots", [("test11", 0.1, 20000), ("surf3", 0.01, 20000),
                                            ("surf5", 0.001, 30000), ("mixed", None, 20000)])
def test_k3_replay_parity(capi, ref, name, p, n_shots):
    """Every sampled shot: outcome from the kernel == reference propagate() of the exported
    fault list; tallies consistent with the packed bits."""
    if name == "test11":
        from qiskit_qec_b200.circuits import EncodedCircuit
        circ = EncodedCircuit(3, 3)
        circ.ops = [[7, 0], [7, 1], [7, 2], [0, 0], [5, 0, 1], [6, 2], [5, 1, 2], [6, 0],
                    [8, 0, 0], [8, 1, 1], [8, 2, 2]]
        idx = [3, 4, 5, 6, 7]
        labels = ["h", "cx", "id", "cx2", "id"]
        lw = {"h": [("x", 1), ("y", 1), ("z", 1)], "cx": [("ix", 1), ("xi", 1), ("xx", 1)],
              "id": [("x", 1), ("y", 1)],
              "cx2": [("ix", 1), ("xi", 1), ("xx", 1), ("iz", 1), ("zi", 1), ("zz", 1)]}
        lp = {k: p for k in lw}
    else:
        circ = {"surf3": surface_code_circuit(3, 3), "surf5": surface_code_circuit(5, 5),
                "mixed": surface_code_circuit(3, 2, basis="x")}[name]
        model = standard_depolarizing_model(p if p is not None else 0.0)
        if name == "mixed":
            for lab, pp in (("h", 0.02), ("cx", 0.005), ("reset", 0.0), ("measure", 0.05)):
                model.set_error_probability(lab, pp)
            model.add_operation("cx", {"ix": 3, "zz": 1, "yy": 0, "xi": 2})
            model.set_error_probability("cx", 0.005)
        idx, labels, lw, lp = sampler_inputs(circ, model)
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    sm, names = make_sampler(capi, g, idx, labels, lw, lp, seed=1234)
    bits, tally, cnt, ev = sm.sample(64, n_shots, want_tally=True, fault_cap=16)
    got = capi.unpack_bits(bits, n_shots)
    per_shot = _decode_faults(cnt, ev, n_shots)
    ep = ref._CErrorPropagator(circ.nq, circ.nc)
    ep.load_circuit(circ.nq, circ.nc, circ.ops)
    nf = np.zeros(64, dtype=np.uint64)
    for s in range(n_shots):
        f = per_shot[s]
        nf[min(len(f), 63)] += 1
        if s % 3 and name == "surf5":
            continue
        want = ep.propagate([idx[l] for l, _ in f], [lw[labels[l]][d][0] for l, d in f])
        assert got[s].tolist() == want, f"shot {s} faults {f}"
    np.testing.assert_array_equal(tally[:circ.nc], got.sum(axis=0).astype(np.uint64))
    assert tally[circ.nc] == np.count_nonzero(got.any(axis=1))
    assert tally[circ.nc + 1] == n_shots
    np.testing.assert_array_equal(tally[circ.nc + 2:], nf)
    # shard independence: same shots drawn in two pieces give the same bits
    b1, _, _, _ = sm.sample(64, 4096)
    b2, _, _, _ = sm.sample(64 + 4096, n_shots - 4096)
    np.testing.assert_array_equal(np.concatenate([b1, b2], axis=1)[:, :bits.shape[1]], bits)


def test_k3_statistics_reference_kats(capi):
    """The two statistical KATs of the reference's test_compiledfaultsampler.py:10-93, at a
    stated binomial 5-sigma instead of the reference's 2 decimal places."""
    from qiskit_qec_b200.circuits import EncodedCircuit
    n = 1 << 20
    circ = EncodedCircuit(1, 1)
    circ.ops = [[7, 0], [0, 0], [8, 0, 0]]
    g = capi.Circuit(1, 1, circ.ops)
    lw = {"h": [("x", 1), ("y", 1), ("z", 1)]}
    sm, _ = make_sampler(capi, g, [1], ["h"], lw, {"h": 0.2}, seed=10)
    _, tally, _, _ = sm.sample(0, n, want_bits=False, want_tally=True)
    pr = 0.2 * 2.0 / 3.0
    assert abs(tally[0] / n - pr) < 5 * np.sqrt(pr * (1 - pr) / n)
    circ = EncodedCircuit(3, 3)
    circ.ops = [[7, 0], [7, 1], [7, 2], [0, 0], [5, 0, 1], [6, 2], [5, 1, 2], [6, 0],
                [8, 0, 0], [8, 1, 1], [8, 2, 2]]
    g = capi.Circuit(3, 3, circ.ops)
    lw = {"h": [("x", 1), ("y", 1), ("z", 1)], "cx": [("ix", 1), ("xi", 1), ("xx", 1)],
          "id": [("x", 1), ("y", 1)],
          "cx2": [("ix", 1), ("xi", 1), ("xx", 1), ("iz", 1), ("zi", 1), ("zz", 1)]}
    sm, _ = make_sampler(capi, g, [3, 4, 5, 6, 7], ["h", "cx", "id", "cx2", "id"], lw,
                         {k: 0.1 for k in lw}, seed=10)
    _, tally, _, _ = sm.sample(0, n, want_bits=False, want_tally=True)
    from math import comb
    for k in range(6):
        pk = comb(5, k) * 0.1 ** k * 0.9 ** (5 - k)
        assert abs(tally[3 + 2 + k] / n - pk) < 5 * np.sqrt(pk * (1 - pk) / n), k
