"""GPU parity on the BASELINE configurations round 1 left without evidence, and on K4:

  * cfg4  SurfaceCodeCircuit d=11 T=11 (nc = 1441, 8801 ops): the sampler that serves circuits
          with more than 254 clbits -- against the CPU restatement of its draw order, against
          explicit propagation (C oracle) and against the compiled reference's propagate(); the
          whole order-1 stream (79 201 paths) against the reference enumerator.
  * cfg5  heavy-hex ARC T=3, order 3: head of the stream against the reference enumerator,
          random / tail windows against the reference propagate() on host-unranked fault lists.
  * K4    the hashed outcome table (qec_outcome_histogram_rows_dev, qec_sampler_counts) against
          numpy's unique() over the packed bits of the same shots.
  * duplicate fault locations on one op: the reference keeps the LAST one
          (errorpropagator.cpp:265-267); the linear paths step aside for the frame kernels.
"""

import numpy as np
import pytest

from helpers import make_enumerator, make_sampler, noise_inputs, ref_outcome_matrix, sampler_inputs
from qiskit_qec_b200.circuits import (EncodedCircuit, arc_circuit, heavy_hex_27_arc_layout,
                                      surface_code_circuit)
from qiskit_qec_b200.noise import standard_depolarizing_model

pytestmark = pytest.mark.gpu


def _decode(cnt, ev, n_shots):
    per_shot = [[] for _ in range(n_shots)]
    for w in range(len(cnt)):
        for e in ev[w, :cnt[w]]:
            loc, bit, dig = int(e) >> 14, (int(e) >> 9) & 31, int(e) & 511
            per_shot[32 * w + bit].append((loc, dig))
    return per_shot


# ------------------------------------------------------------------------------------ cfg4
@pytest.mark.parametrize("d,T,p,n_shots,lo", [(11, 3, 2e-3, 6000, 8192 * 2 + 32 * 7),
                                               (11, 11, 1e-3, 5000, 0),
                                               (11, 11, 1e-3, 2100, 512 * 3 + 96)])
def test_cfg4_sampler_vs_port_oracle_and_reference(capi, port, ref, d, T, p, n_shots, lo):
    """d=11: fault lists == the CPU restatement of the kernel's draw order; outcomes == explicit
    propagation of those lists by the C oracle (every shot) and by the reference's propagate()
    (2 000+ shots of the T=11 circuit); tallies == the bits.  The windows start off a batch
    boundary and end on a ragged tail."""
    circ = surface_code_circuit(d, T)
    model = standard_depolarizing_model(p)
    idx, labels, lw, lp = sampler_inputs(circ, model)
    _, _, paulis = noise_inputs(circ, model)
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    sm, _ = make_sampler(capi, g, idx, labels, lw, lp, seed=2024)
    bits, tally, cnt, ev = sm.sample(lo, n_shots, want_tally=True, fault_cap=512)
    kernel = capi.last_kernel()
    assert "sample:linear" in kernel and circ.nc > 254, kernel
    tab = sm.tables()
    nw = (n_shots + 31) // 32
    paths = [([], []) for _ in range(n_shots)]
    for w in range(nw):
        nv = min(32, n_shots - 32 * w)
        want_ev = port.sample_word(tab, lo // 32 + w, nv)
        assert np.array_equal(ev[w, :cnt[w]], want_ev), f"fault list of word {w} differs from the port"
        for e in want_ev:
            loc, bit, dig = int(e) >> 14, (int(e) >> 9) & 31, int(e) & 511
            paths[32 * w + bit][0].append(idx[loc])
            paths[32 * w + bit][1].append(paulis[loc][dig])
    want = port.propagate_batch(circ.nq, circ.nc, circ.ops, paths)
    got = capi.unpack_bits(bits, n_shots)
    assert np.array_equal(got, want), "d=11 outcomes differ from explicit propagation"
    nc = circ.nc
    assert np.array_equal(tally[:nc], want.sum(axis=0).astype(np.uint64))
    assert int(tally[nc]) == int(want.any(axis=1).sum()) and int(tally[nc + 1]) == n_shots
    k = np.minimum(np.array([len(pp[0]) for pp in paths]), 63)
    assert np.array_equal(tally[nc + 2:nc + 66], np.bincount(k, minlength=64).astype(np.uint64))
    # mean faults per shot is what the noise model says (8.8 at T=11, p=1e-3), 5 sigma
    mean = len(idx) * p
    assert abs(k.mean() - mean) < 5 * np.sqrt(mean / n_shots)
    # the compiled reference on the same fault lists
    ep = ref._CErrorPropagator(circ.nq, circ.nc)
    ep.load_circuit(circ.nq, circ.nc, circ.ops)
    step = 1 if T == 11 and n_shots <= 2100 else 3
    for s in range(0, n_shots, step):
        assert got[s].tolist() == ep.propagate(*paths[s]), f"shot {s}"
    # the bulk call (no export) and a tallies-only call agree with it
    bits2, tally2, _, _ = sm.sample(lo, n_shots, want_tally=True)
    assert np.array_equal(bits2, bits) and np.array_equal(tally2, tally)
    _, tally3, _, _ = sm.sample(lo, n_shots, want_bits=False, want_tally=True)
    assert np.array_equal(tally3, tally)


def test_cfg4_order1_full_stream_vs_reference(capi, ref):
    """All 79 201 order-1 fault paths of d=11 T=11 against reference _CFaultEnumerator
    (SURVEY 8(c)(1)): index, labels, errors via host unranking; every outcome bit."""
    circ = surface_code_circuit(11, 11)
    idx, labels, paulis = noise_inputs(circ, standard_depolarizing_model())
    fe = ref._CFaultEnumerator(1, circ.nq, circ.nc, circ.ops, idx, labels, paulis)
    ref_paths = []
    while not fe.done():
        ref_paths += fe.enumerate(20000)
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    en = make_enumerator(capi, g, 1, idx, paulis)
    n = en.num_paths()
    assert n == len(ref_paths) == 79201
    bits, tally = en.range(0, n, want_tally=True)
    got = capi.unpack_bits(bits, n)
    want = ref_outcome_matrix(ref_paths)
    assert np.array_equal(got, want)
    assert np.array_equal(tally[:circ.nc], want.sum(axis=0).astype(np.uint64))
    for r in range(0, n, 997):
        locs, digs = en.unrank(r)
        assert ref_paths[r][0] == r and ref_paths[r][1] == [labels[locs[0]]]
        assert ref_paths[r][2] == [paulis[locs[0]][digs[0]]]


# ------------------------------------------------------------------------------------ cfg5
def test_cfg5_arc_order3_vs_reference(capi, ref):
    """BASELINE cfg5 (heavy-hex ARC, T=3, order 3, 3.4e8 paths): the first 100 000 paths against
    the reference stream (index, labels, errors, outcome), then five random / tail windows against
    reference propagate() on host-unranked fault lists."""
    links, sched, color = heavy_hex_27_arc_layout()
    circ = arc_circuit(links, 3, "zx", sched, color)
    idx, labels, paulis = noise_inputs(circ, standard_depolarizing_model())
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    en = make_enumerator(capi, g, 3, idx, paulis)
    n = en.num_paths()
    assert n > 3 * 10**8
    fe = ref._CFaultEnumerator(3, circ.nq, circ.nc, circ.ops, idx, labels, paulis)
    head = []
    while len(head) < 100_000:
        head += fe.enumerate(20000)
    m = 100_000
    bits, tally = en.range(0, m, want_tally=True)
    assert "enumerate:linear" in capi.last_kernel()
    got = capi.unpack_bits(bits, m)
    want = ref_outcome_matrix(head[:m])
    assert np.array_equal(got, want)
    assert np.array_equal(tally[:circ.nc], want.sum(axis=0).astype(np.uint64))
    assert int(tally[circ.nc]) == int(want.any(axis=1).sum())
    for r in (0, 1, 31, 32, 4097, 65535, m - 1):
        locs, digs = en.unrank(r)
        assert head[r][0] == r
        assert head[r][1] == [labels[l] for l in locs]
        assert head[r][2] == [paulis[l][dd] for l, dd in zip(locs, digs)]
    ep = ref._CErrorPropagator(circ.nq, circ.nc)
    ep.load_circuit(circ.nq, circ.nc, circ.ops)
    rng = np.random.default_rng(17)
    for lo in [n - 2500] + [int(x) for x in rng.integers(0, n - 2500, size=4)]:
        hi = lo + 2500 - int(rng.integers(0, 40))
        wbits, _ = en.range(lo, hi)
        wgot = capi.unpack_bits(wbits, hi - lo)
        for r in range(lo, hi, 5):
            locs, digs = en.unrank(r)
            want_r = ep.propagate([idx[l] for l in locs], [paulis[l][dd] for l, dd in zip(locs, digs)])
            assert wgot[r - lo].tolist() == want_r, f"path {r}"


# -------------------------------------------------------------------------------------- K4
def _np_counts(packed, n):
    rows = capi_pack_rows(packed, n)
    nz = rows.any(axis=1)
    uniq, cnt = np.unique(rows[nz], axis=0, return_counts=True)
    return uniq, cnt, int((~nz).sum())


def capi_pack_rows(packed, n):
    from qiskit_qec_b200 import _capi
    return _capi.pack_rows(packed, n)


@pytest.mark.parametrize("d,T,p,n_shots,lo", [(3, 3, 0.01, 50_000, 0), (5, 5, 1e-3, 300_000, 4096 * 3),
                                               (5, 5, 5e-3, 70_001, 64), (11, 3, 1e-3, 20_000, 0)])
def test_k4_sampler_counts_vs_numpy_unique(capi, d, T, p, n_shots, lo):
    """qec_sampler_counts (sample + hashed table on the device, only distinct outcomes copied
    back) == numpy unique() over the packed bits of the same shots: same set of outcomes, same
    multiplicities, same all-zero count; table keys == the documented 64-bit hash."""
    circ = surface_code_circuit(d, T)
    idx, labels, lw, lp = sampler_inputs(circ, standard_depolarizing_model(p))
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    sm, _ = make_sampler(capi, g, idx, labels, lw, lp, seed=77)
    bits, tally, _, _ = sm.sample(lo, n_shots, want_tally=True)
    uniq, cnt, n_zero = _np_counts(bits, n_shots)
    rows, counts, got_zero, tally2 = sm.counts(lo, n_shots, want_tally=True)
    assert got_zero == n_zero and int(counts.sum()) + got_zero == n_shots
    order = np.lexsort(rows.T[::-1])
    rows, counts = rows[order], counts[order]
    order_u = np.lexsort(uniq.T[::-1])
    assert np.array_equal(rows, uniq[order_u]) and np.array_equal(counts, cnt[order_u].astype(np.uint64))
    assert np.array_equal(tally2, tally)
    assert len(np.unique(capi.outcome_hash(rows))) == len(rows)   # no two outcomes share a key
    # a table that is too small is reported, never silently wrong
    if len(rows) > 64:
        with pytest.raises(RuntimeError, match="outcome table full"):
            sm.counts(lo, n_shots, table_cap=32)


def test_k4_histogram_dev_entry_point(capi):
    """The device-pointer entry point over an enumerator's packed outcomes (order 2, d=3): table
    (keys, counts, rows, status) == numpy."""
    import torch
    circ = surface_code_circuit(3, 2)
    idx, labels, paulis = noise_inputs(circ, standard_depolarizing_model())
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    en = make_enumerator(capi, g, 2, idx, paulis)
    n = en.num_paths()
    nw = (n + 31) // 32
    dev = torch.device("cuda", 0)
    dbits = torch.zeros((circ.nc, nw), dtype=torch.int32, device=dev)
    en.range_dev(0, n, dbits.data_ptr(), 0)
    cap = 1 << 16
    rw = (circ.nc + 31) // 32
    keys = torch.zeros(cap, dtype=torch.int64, device=dev)
    counts = torch.zeros(cap, dtype=torch.int64, device=dev)
    rows = torch.zeros((cap, rw), dtype=torch.int32, device=dev)
    status = torch.zeros(3, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    capi.outcome_histogram_dev(g, dbits.data_ptr(), n, keys.data_ptr(), counts.data_ptr(), cap,
                               d_rows=rows.data_ptr(), d_status=status.data_ptr())
    torch.cuda.synchronize()
    packed = dbits.cpu().numpy().view(np.uint32)
    uniq, cnt, n_zero = _np_counts(packed, n)
    k = keys.cpu().numpy().view(np.uint64)
    occ = k != 0
    st = status.cpu().numpy()
    assert int(st[0]) == int(occ.sum()) == len(uniq) and int(st[1]) == 0 and int(st[2]) == n_zero
    got_rows = rows.cpu().numpy().view(np.uint32)[occ]
    got_cnt = counts.cpu().numpy()[occ]
    assert np.array_equal(capi.outcome_hash(got_rows), k[occ])
    o1, o2 = np.lexsort(got_rows.T[::-1]), np.lexsort(uniq.T[::-1])
    assert np.array_equal(got_rows[o1], uniq[o2]) and np.array_equal(got_cnt[o1], cnt[o2])


def test_sample_counts_driver_uses_the_device_table(capi):
    """FaultSampler.sample_counts (the Qiskit-counts adaptor) is served by K4: same dictionary
    as counting the packed bits of the same shots on the host."""
    from qiskit_qec_b200.analysis.faultsampler import FaultSampler, counts_from_packed
    circ = surface_code_circuit(3, 2)
    model = standard_depolarizing_model(0.02)
    fs = FaultSampler(circ, model, sim_seed=5)
    n = 40_000
    counts = fs.sample_counts(n)
    fs.faultsamp.set_cursor(0)
    packed, _ = fs.sample_packed(n)
    assert counts == counts_from_packed(np.asarray(packed), n, fs.reg_sizes)
    assert sum(counts.values()) == n


# ------------------------------------------------------------------------ duplicate locations
def test_duplicate_location_ops_keep_the_reference_rule(capi, ref):
    """Two fault locations on ONE op: the reference's propagate() keeps only the last fault of
    an op.  The linear (XOR of signatures) paths must not serve such inputs."""
    circ = EncodedCircuit(3, 3)
    circ.ops = [[0, 0], [5, 0, 1], [5, 1, 2], [0, 2], [8, 0, 0], [8, 1, 1], [8, 2, 2]]
    idx = [0, 1, 1, 2, 3]
    labels = ["h", "cx", "cx", "cx", "h"]
    paulis = [["x", "y", "z"], ["ix", "xi", "zz"], ["xx", "iz"], ["ix", "zi"], ["x", "z"]]
    for order in (1, 2, 3):
        fe = ref._CFaultEnumerator(order, 3, 3, circ.ops, idx, labels, paulis)
        want = []
        while not fe.done():
            want += fe.enumerate(1000)
        g = capi.Circuit(3, 3, circ.ops)
        en = make_enumerator(capi, g, order, idx, paulis)
        n = en.num_paths()
        assert n == len(want)
        bits, _ = en.range(0, n)
        assert "linear" not in capi.last_kernel(), capi.last_kernel()
        assert np.array_equal(capi.unpack_bits(bits, n), ref_outcome_matrix(want))
    lw = {"h": [("x", 1), ("z", 1)], "cx": [("ix", 1), ("xi", 1), ("zz", 1), ("xx", 2)]}
    g = capi.Circuit(3, 3, circ.ops)
    sm, _ = make_sampler(capi, g, idx, labels, lw, {"h": 0.3, "cx": 0.4}, seed=3)
    n = 4096
    bits, _, cnt, ev = sm.sample(0, n, fault_cap=32)
    assert "linear" not in capi.last_kernel(), capi.last_kernel()
    got = capi.unpack_bits(bits, n)
    ep = ref._CErrorPropagator(3, 3)
    ep.load_circuit(3, 3, circ.ops)
    for s, f in enumerate(_decode(cnt, ev, n)):
        want = ep.propagate([idx[l] for l, _ in f], [lw[labels[l]][dd][0] for l, dd in f])
        assert got[s].tolist() == want, f"shot {s} faults {f}"


# ------------------------------------------------------------------- compact output (events)
@pytest.mark.parametrize("d,T,p,n_shots,lo,mode", [
    (3, 3, 0.01, 50_001, 0, None), (5, 5, 1e-3, 700_000, 4096 * 3 + 64, None),
    (5, 5, 2e-3, 100_000, 32, "lanes16"), (5, 5, 2e-3, 40_000, 0, "warp"), (11, 3, 1e-3, 20_000, 0, None)])
def test_compact_events_equal_the_fault_export(capi, monkeypatch, d, T, p, n_shots, lo, mode):
    """qec_sampler_sample_events (fault lists compacted on the device, ~3 B/shot over PCIe) holds
    exactly the faults of the dense call's export, word by word; tallies are the dense call's; and
    the outcome of every shot is the XOR of its faults' order-1 outcomes (what makes the compact
    form lossless), checked against the dense bits."""
    if mode:
        monkeypatch.setenv("QEC_B200_LIN_MODE", mode)
    circ = surface_code_circuit(d, T)
    model = standard_depolarizing_model(p)
    idx, labels, lw, lp = sampler_inputs(circ, model)
    _, _, paulis = noise_inputs(circ, model)
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    sm, _ = make_sampler(capi, g, idx, labels, lw, lp, seed=31)
    n_small = min(n_shots, 60_000)
    bits, tally, cnt, ev = sm.sample(lo, n_small, want_tally=True, fault_cap=8)
    wc, events, tally_c = sm.sample_events(lo, n_small, want_tally=True)
    assert np.array_equal(tally_c, tally)
    assert np.array_equal(wc.astype(np.uint32), cnt)
    wc3, ev3, _ = sm.sample_events(lo, n_small, packed24=True)   # the 24-bit form, whichever kernel exports
    e3 = capi.unpack_events24(ev3)
    assert np.array_equal(wc3, wc) and len(e3) == len(events)
    word_of = np.repeat(np.arange(len(wc)), wc.astype(np.int64))  # (the CTA variant appends with atomics:
    key = lambda e: np.lexsort((e, word_of))                      #  a word's events come in any order)
    assert np.array_equal(e3[key(e3)], events[key(events)])
    off = np.concatenate([[0], np.cumsum(wc.astype(np.int64))])
    assert off[-1] == len(events)
    first = np.cumsum([0] + [len(pl) for pl in paulis])
    en = make_enumerator(capi, g, 1, idx, paulis)
    sig = capi.unpack_bits(en.range(0, en.num_paths())[0], en.num_paths())   # order-1 outcomes
    got = capi.unpack_bits(bits, n_small)
    want = np.zeros_like(got)
    for w in range(len(wc)):
        e = np.sort(events[off[w]:off[w + 1]])
        assert np.array_equal(e, np.sort(ev[w, :cnt[w]])), f"word {w}"
        for x in e:
            loc, bit, dig = int(x) >> 14, (int(x) >> 9) & 31, int(x) & 511
            want[32 * w + bit] ^= sig[first[loc] + dig]
    assert np.array_equal(got, want)
    # the full size: counts add up, tallies agree with a tallies-only dense call
    wc, events, tally_c = sm.sample_events(lo, n_shots, want_tally=True)
    _, tally_d, _, _ = sm.sample(lo, n_shots, want_bits=False, want_tally=True)
    assert int(wc.astype(np.int64).sum()) == len(events) and np.array_equal(tally_c, tally_d)
    k = np.arange(64, dtype=np.uint64)
    assert int((tally_c[circ.nc + 2:] * k).sum()) == len(events) or int(tally_c[circ.nc + 2 + 63]) > 0


# ------------------------------------------------------------------------- 8(f2): heavy-hexagon
@pytest.mark.parametrize("kw", [dict(), dict(num_initialize=1, idle_before_measure=True),
                                dict(num_initialize=0, group_meas=True, blocks=2)])
def test_hhc_labelled_idles_order1_stream_vs_reference(capi, ref, kw):
    """HHC circuit with its labelled idles as separate noise channels (idm / idbm<k> / idinit /
    plain id): the whole order-1 stream -- index, labels, errors, outcomes -- against the reference
    enumerator; an order-2 window too."""
    from qiskit_qec_b200.hhc import hhc_circuit
    from qiskit_qec_b200.noise import PauliNoiseModel
    circ = hhc_circuit(d=3, **kw)
    model = PauliNoiseModel()
    model.add_operation("cx", {p: 1 for p in ["ix", "iy", "iz", "xi", "xx", "xy", "xz", "yi", "yx", "yy", "yz",
                                               "zi", "zx", "zy", "zz"]})
    model.add_operation("h", {"x": 1, "y": 1, "z": 1})
    model.add_operation("id", {"x": 1, "y": 1, "z": 1})
    model.add_operation("idm", {"x": 1, "y": 2, "z": 1})
    model.add_operation("idinit", {"x": 1})
    model.add_operation("reset", {"x": 1})
    model.add_operation("measure", {"x": 1})
    for lab in sorted(set(circ.labels)):
        if lab.startswith("idbm"):
            model.add_operation(lab, {"x": 1, "z": 1})
    idx, labels, paulis = noise_inputs(circ, model)
    assert "idm" in labels or kw.get("group_meas")
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    for order, limit in ((1, None), (2, 30_000)):
        fe = ref._CFaultEnumerator(order, circ.nq, circ.nc, circ.ops, idx, labels, paulis)
        want = []
        while not fe.done() and (limit is None or len(want) < limit):
            want += fe.enumerate(10_000)
        en = make_enumerator(capi, g, order, idx, paulis)
        n = len(want) if limit else en.num_paths()
        assert n == len(want)
        bits, _ = en.range(0, n)
        assert np.array_equal(capi.unpack_bits(bits, n), ref_outcome_matrix(want))
        for r in range(0, n, max(1, n // 200)):
            locs, digs = en.unrank(r)
            assert want[r][0] == r and want[r][1] == [labels[l] for l in locs]
            assert want[r][2] == [paulis[l][dd] for l, dd in zip(locs, digs)]


def test_events24_equal_the_32_bit_events(capi):
    """qec_sampler_sample_events24 (three bytes per fault on the wire) decodes to exactly the events
    of qec_sampler_sample_events, counts and tallies included; circuits it cannot encode are refused."""
    from qiskit_qec_b200.circuits import surface_code_circuit
    from qiskit_qec_b200.noise import standard_depolarizing_model
    from helpers import make_sampler, sampler_inputs
    for d, T, n, lo in ((5, 5, 300_000 + 17, 64), (3, 2, 10_000, 0), (7, 3, 100_000 + 5, 4096)):
        circ = surface_code_circuit(d, T)
        idx, labels, lw, lp = sampler_inputs(circ, standard_depolarizing_model(2e-3))
        g = capi.Circuit(circ.nq, circ.nc, circ.ops)
        sm, _ = make_sampler(capi, g, idx, labels, lw, lp, seed=21)
        wc, ev, tl = sm.sample_events(lo, n, want_tally=True)
        wc3, ev3, tl3 = sm.sample_events(lo, n, want_tally=True, packed24=True)
        assert ev3.dtype == np.uint8 and ev3.shape == (len(ev), 3)
        assert np.array_equal(wc, wc3) and np.array_equal(tl, tl3)
        # (inside a word the order is the kernel's; the CTA variant appends with atomics)
        word_of = np.repeat(np.arange(len(wc)), wc.astype(np.int64))
        e3 = capi.unpack_events24(ev3)
        assert np.array_equal(e3[np.lexsort((e3, word_of))], ev[np.lexsort((ev, word_of))])


def test_events24_refuses_what_it_cannot_encode(capi):
    """More than 16 Paulis on a location do not fit the 4-bit digit of the 24-bit event."""
    circ = EncodedCircuit(nq=2, nc=2, name="wide")
    circ.cx(0, 1)
    circ.measure(0, 0)
    circ.measure(1, 1)
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    two = ["ix", "iy", "iz", "xi", "xx", "xy", "xz", "yi", "yx", "yy", "yz", "zi", "zx", "zy", "zz", "xx", "zz"]
    sm = capi.Sampler(g, [0], [0], [0.2], [[capi.pauli_code(p) for p in two]], [[1.0] * len(two)], 3)
    wc, ev, _ = sm.sample_events(0, 4096)
    assert int(wc.sum()) == len(ev) > 0
    with pytest.raises(RuntimeError, match="24-bit events"):
        sm.sample_events(0, 4096, packed24=True)


def test_events_do_not_depend_on_the_chunking(capi):
    """qec_sampler_sample_events streams its output in chunks (a short first one, then equal ones);
    a call that spans several chunks returns exactly the concatenation of smaller calls."""
    from qiskit_qec_b200.circuits import surface_code_circuit
    from qiskit_qec_b200.noise import standard_depolarizing_model
    from helpers import make_sampler, sampler_inputs
    circ = surface_code_circuit(3, 2)
    idx, labels, lw, lp = sampler_inputs(circ, standard_depolarizing_model(0.01))
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    sm, _ = make_sampler(capi, g, idx, labels, lw, lp, seed=4)
    n, cut = 6_000_000 + 32 * 5 + 9, 2_000_000 + 1024 * 3 + 64
    for packed in (False, True):
        wc, ev, tl = sm.sample_events(0, n, want_tally=True, packed24=packed)
        wa, ea, ta = sm.sample_events(0, cut, want_tally=True, packed24=packed)
        wb, eb, tb = sm.sample_events(cut, n - cut, want_tally=True, packed24=packed)
        assert np.array_equal(wc, np.concatenate([wa, wb]))
        assert np.array_equal(ev, np.concatenate([ea, eb]))
        assert np.array_equal(tl, ta + tb)


def test_events16_delta_coded_form(capi, monkeypatch):
    """qec_sampler_sample_events16 (two bytes per entry, delta-coded in slot order) decodes to the
    events of the 32-bit form, word by word; served by the lane-stream kernel only."""
    from qiskit_qec_b200.circuits import surface_code_circuit
    from qiskit_qec_b200.noise import standard_depolarizing_model
    from helpers import make_sampler, sampler_inputs
    for d, T, p, n, lo in ((5, 5, 1e-3, 400_000 + 21, 4096 + 64), (3, 2, 0.02, 20_000, 0), (5, 5, 2e-5, 300_000, 32)):
        circ = surface_code_circuit(d, T)
        idx, labels, lw, lp = sampler_inputs(circ, standard_depolarizing_model(p))
        g = capi.Circuit(circ.nq, circ.nc, circ.ops)
        sm, _ = make_sampler(capi, g, idx, labels, lw, lp, seed=23)
        wc, ev, tl = sm.sample_events(lo, n, want_tally=True)
        assert "lane-streams" in capi.last_kernel()
        wc2, ent, tl2 = sm.sample_events(lo, n, want_tally=True, packed16=True)
        assert ent.dtype == np.uint16 and np.array_equal(tl, tl2) and int(wc2.sum()) == len(ent)
        cnt, dec = capi.unpack_events16(wc2, ent, sm.tables())
        assert np.array_equal(cnt, wc)
        word_of = np.repeat(np.arange(len(wc)), wc.astype(np.int64))
        assert np.array_equal(dec[np.lexsort((dec, word_of))], ev[np.lexsort((ev, word_of))])
    # the CTA variant appends a word's events with atomics, in no order: refused
    monkeypatch.setenv("QEC_B200_LIN_MODE", "cta")
    circ = surface_code_circuit(3, 2)
    idx, labels, lw, lp = sampler_inputs(circ, standard_depolarizing_model(0.01))
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    sm, _ = make_sampler(capi, g, idx, labels, lw, lp, seed=23)
    with pytest.raises(RuntimeError, match="16-bit events"):
        sm.sample_events(0, 4096, packed16=True)
