"""GPU parity of the linear-response sampler (qec_linear.cu).

The kernel never walks the gate list: it XORs per-fault outcome signatures.  These tests pin
that shortcut to explicit propagation:
  * the premise -- every order-2 outcome of the frame kernels is the XOR of two order-1
    outcomes -- over whole enumerations;
  * every sampled shot: packed outcome == explicit propagation (CPU oracle) of the shot's
    exported fault list, and the fault lists == the CPU restatement of the draw order;
  * window / shard invariance, tallies vs bits, thinning with unequal class probabilities,
    tiny p (first-level gap table exhausted), p close to 1;
  * fault rates against the closed form, 5 sigma.
test_gpu_parity.py::test_k3_replay_parity additionally replays the exported faults through the
compiled reference's propagate()."""

import numpy as np
import pytest

from qiskit_qec_b200.circuits import (arc_circuit, heavy_hex_27_arc_layout, repetition_code_circuit,
                                      surface_code_circuit)
from qiskit_qec_b200.noise import standard_depolarizing_model

from helpers import make_enumerator, make_sampler, noise_inputs, sampler_inputs

pytestmark = pytest.mark.gpu


def _circuit(name):
    if name == "rep":
        return repetition_code_circuit(3, 2)
    if name == "surf3":
        return surface_code_circuit(3, 3)
    if name == "surf5":
        return surface_code_circuit(5, 5)
    if name == "surf3x":
        return surface_code_circuit(3, 2, basis="x")
    if name == "surf7":
        return surface_code_circuit(7, 3)
    if name == "rep100":  # two chains of 100 rows: the lane read-out program outgrows its unrolled trips
        return repetition_code_circuit(3, 100, resets=True)
    links, sched, color = heavy_hex_27_arc_layout()
    return arc_circuit(links, 3, "zx", sched, color)


@pytest.mark.parametrize("name", ["rep", "surf3", "surf3x", "arc"])
def test_order2_outcomes_are_xors_of_order1_outcomes(capi, name):
    """GF(2) linearity of the frame kernels, path by path, over a whole order-2 enumeration
    (or its first 300 000 paths)."""
    circ = _circuit(name)
    idx, labels, paulis = noise_inputs(circ, standard_depolarizing_model())
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    e1 = make_enumerator(capi, g, 1, idx, paulis)
    n1 = e1.num_paths()
    sig = capi.unpack_bits(e1.range(0, n1)[0], n1)  # [fault][clbit]
    first = np.cumsum([0] + [len(p) for p in paulis])
    e2 = make_enumerator(capi, g, 2, idx, paulis)
    n2 = min(e2.num_paths(), 300_000)
    got = capi.unpack_bits(e2.range(0, n2)[0], n2)
    step = max(1, n2 // 4000)
    for r in range(0, n2, step):
        locs, digs = e2.unrank(r)
        want = sig[first[locs[0]] + digs[0]] ^ sig[first[locs[1]] + digs[1]]
        assert np.array_equal(got[r], want), f"path {r}"


@pytest.mark.parametrize("name,p,n_shots,lo", [
    ("rep", 0.05, 20_000, 0), ("surf3", 0.01, 30_000, 4096 * 3 + 32 * 5), ("surf3x", 0.3, 6_000, 32),
    ("surf5", 1e-3, 40_000, 8192 - 64), ("arc", 2e-3, 20_000, 0), ("surf7", 5e-3, 9_000, 0)])
def test_linear_sampler_equals_explicit_propagation(capi, port, name, p, n_shots, lo):
    _check_sampler_against_port_and_oracle(capi, port, name, p, n_shots, lo)


@pytest.mark.parametrize("mode,draw,kernel", [
    ("lanes16", None, "lane16-streams"), ("lanes", None, "lane-streams"), ("warp", None, "warp-batches"),
    ("warp", "table", "draw=0"),
    ("warp", "computed1", "draw=1"), ("cta", None, "cta-batches"), ("cta", "table", "draw=0")])
@pytest.mark.parametrize("name,p,n_shots,lo", [("surf3", 0.01, 20_000, 4096 + 32 * 3), ("rep", 0.05, 9_000, 0),
                                               ("surf5", 2e-3, 12_000, 1024 * 5 - 96)])
def test_linear_sampler_every_kernel_and_draw_scheme(capi, port, monkeypatch, mode, draw, kernel, name, p, n_shots, lo):
    """The four kernels (16- and 32-shot lane streams / warp batches / CTA batches) and the three draw schemes
    (gap tables, computed gaps with one or two candidates per Philox call), each against the CPU
    restatement of its draw order and against explicit propagation."""
    monkeypatch.setenv("QEC_B200_LIN_MODE", mode)
    if draw:
        monkeypatch.setenv("QEC_B200_LIN_DRAW", draw)
    _check_sampler_against_port_and_oracle(capi, port, name, p, n_shots, lo)
    assert kernel in capi.last_kernel(), capi.last_kernel()


def _check_sampler_against_port_and_oracle(capi, port, name, p, n_shots, lo):
    circ = _circuit(name)
    model = standard_depolarizing_model(p)
    idx, labels, lw, lp = sampler_inputs(circ, model)
    _, _, paulis = noise_inputs(circ, model)
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    sm, _ = make_sampler(capi, g, idx, labels, lw, lp, seed=99)
    bits, tally, cnt, ev = sm.sample(lo, n_shots, want_tally=True, fault_cap=64)
    assert "sample:linear" in capi.last_kernel()
    tab = sm.tables()
    assert int(tab[7]) & 2
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
    assert np.array_equal(got, want), "linear-response outcomes differ from explicit propagation"
    # tallies: per-clbit ones, shots with any 1, shot count, faults-per-shot histogram
    nc = circ.nc
    assert np.array_equal(tally[:nc], want.sum(axis=0).astype(np.uint64))
    assert int(tally[nc]) == int(want.any(axis=1).sum()) and int(tally[nc + 1]) == n_shots
    k = np.minimum(np.array([len(pp[0]) for pp in paths]), 63)
    assert np.array_equal(tally[nc + 2:nc + 66], np.bincount(k, minlength=64).astype(np.uint64))
    # the bulk call (no export) returns the same bits
    bits2, tally2, _, _ = sm.sample(lo, n_shots, want_tally=True)
    assert np.array_equal(bits2, bits) and np.array_equal(tally2, tally)


def test_linear_sampler_windows_and_shards(capi):
    """Any 32-aligned window of the shot stream, cut anywhere relative to the kernel's batches,
    returns the bits of the one big call; shard tallies add up."""
    circ = _circuit("surf5")
    idx, labels, lw, lp = sampler_inputs(circ, standard_depolarizing_model(2e-3))
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    sm, _ = make_sampler(capi, g, idx, labels, lw, lp, seed=5)
    n = 3 * 4096 * 7 + 32 * 11 + 5
    bits, tally, _, _ = sm.sample(0, n, want_tally=True)
    full = capi.unpack_bits(bits, n)
    rng = np.random.default_rng(0)
    for _ in range(6):
        lo = int(rng.integers(0, n // 32 - 8)) * 32
        m = int(rng.integers(1, n - lo + 1))
        b, _, _, _ = sm.sample(lo, m)
        assert np.array_equal(capi.unpack_bits(b, m), full[lo:lo + m]), (lo, m)
    cuts = [0, 32 * 100, 4096 * 5 + 64, 4096 * 11, n]
    acc = np.zeros_like(tally)
    for a, b in zip(cuts[:-1], cuts[1:]):
        _, t, _, _ = sm.sample(a, b - a, want_bits=False, want_tally=True)
        acc += t
    assert np.array_equal(acc, tally)


def test_linear_sampler_class_thinning_and_rates(capi):
    """Unequal class probabilities (thinning at pmax), a zero-probability class and non-uniform
    Pauli weights: per-class fault counts and per-Pauli shares within 5 sigma."""
    circ = _circuit("surf3")
    model = standard_depolarizing_model(0.0)
    for lab, pp in (("h", 0.02), ("cx", 0.004), ("reset", 0.0), ("measure", 0.03), ("id", 0.001)):
        model.set_error_probability(lab, pp)
    model.add_operation("cx", {"ix": 3, "zz": 1, "yy": 0, "xi": 2})
    model.set_error_probability("cx", 0.004)
    idx, labels, lw, lp = sampler_inputs(circ, model)
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    sm, names = make_sampler(capi, g, idx, labels, lw, lp, seed=11)
    n = 400_000
    _, _, cnt, ev = sm.sample(0, n, want_bits=False, fault_cap=96)
    assert "sample:linear" in capi.last_kernel()
    evs = np.concatenate([ev[w, :cnt[w]] for w in range(len(cnt))])
    loc, dig = evs >> 14, evs & 511
    lab_of = np.array([names.index(l) for l in labels])
    for k, name in enumerate(names):
        sel = lab_of[loc] == k
        n_loc = int((lab_of == k).sum())
        p = lp.get(name, 0.0)
        mean, sd = n * n_loc * p, np.sqrt(n * n_loc * p * (1 - p))
        assert abs(int(sel.sum()) - mean) <= 5 * sd + 1e-9, (name, int(sel.sum()), mean)
        if p > 0:
            ws = np.array([w for _, w in lw[name]], dtype=float)
            share = np.bincount(dig[sel], minlength=len(ws)) / max(1, int(sel.sum()))
            want = ws / ws.sum()
            sd_s = np.sqrt(want * (1 - want) / max(1, int(sel.sum())))
            assert np.all(np.abs(share - want) <= 5 * sd_s + 1e-12), (name, share, want)


def test_linear_sampler_tiny_and_huge_p(capi, port):
    """p = 1e-7: the first-level gap table is exhausted by most draws (advance K1 * B slots,
    no fault); p = 0.9: a fault on nearly every slot.  Both match the port and explicit
    propagation."""
    circ = _circuit("surf3")
    for p, n in ((1e-7, 2_000_000), (0.9, 2_048)):
        model = standard_depolarizing_model(p)
        idx, labels, lw, lp = sampler_inputs(circ, model)
        _, _, paulis = noise_inputs(circ, model)
        g = capi.Circuit(circ.nq, circ.nc, circ.ops)
        sm, _ = make_sampler(capi, g, idx, labels, lw, lp, seed=3)
        cap = 16 if p < 0.5 else 32 * len(idx) + 8
        bits, tally, cnt, ev = sm.sample(0, n, want_tally=True, fault_cap=cap)
        assert "sample:linear" in capi.last_kernel()
        tab = sm.tables()
        total = int(cnt.sum())
        mean = n * len(idx) * p
        assert abs(total - mean) <= 5 * np.sqrt(mean) + 5
        got = capi.unpack_bits(bits, n)
        words = np.nonzero(cnt)[0][:200] if p < 0.5 else np.arange(len(cnt))
        paths, rows = [], []
        for w in words:
            want_ev = port.sample_word(tab, int(w), min(32, n - 32 * int(w)))
            assert np.array_equal(ev[w, :cnt[w]], want_ev)
            per = {}
            for e in want_ev:
                loc, bit, dig = int(e) >> 14, (int(e) >> 9) & 31, int(e) & 511
                per.setdefault(bit, ([], []))
                per[bit][0].append(idx[loc])
                per[bit][1].append(paulis[loc][dig])
            for bit, pth in per.items():
                paths.append(pth)
                rows.append(32 * int(w) + bit)
        want = port.propagate_batch(circ.nq, circ.nc, circ.ops, paths)
        assert np.array_equal(got[rows], want)
        if p < 0.5:  # every shot without a fault reads all zeros
            mask = np.ones(n, dtype=bool)
            for w in np.nonzero(cnt)[0]:
                mask[32 * w:32 * w + 32] = False
            assert not got[mask].any()


def test_frame_mode_switch(capi, monkeypatch):
    """QEC_B200_SAMPLER=frame keeps the frame-propagation samplers (their own draw order)."""
    circ = _circuit("surf3")
    idx, labels, lw, lp = sampler_inputs(circ, standard_depolarizing_model(0.01))
    monkeypatch.setenv("QEC_B200_SAMPLER", "frame")
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    sm, _ = make_sampler(capi, g, idx, labels, lw, lp, seed=1)
    sm.sample(0, 4096)
    assert "sample:linear" not in capi.last_kernel()
    assert not int(sm.tables()[7]) & 2


@pytest.mark.parametrize("name,order", [("rep", 1), ("rep", 2), ("rep", 3), ("surf3", 1), ("surf3", 2),
                                        ("surf3x", 2), ("arc", 2), ("arc", 3), ("surf5", 1), ("surf5", 2),
                                        ("rep100", 1), ("rep100", 2)])
def test_linear_enumerator_equals_frame_enumerator(capi, monkeypatch, name, order):
    """k_enum_lanes (outcome = XOR of the faults' signatures) against the frame-propagation
    enumerators, bit for bit: whole enumerations when small, else windows cut anywhere (unaligned
    starts, ragged ends, the last paths), bits and tallies."""
    circ = _circuit(name)
    idx, labels, paulis = noise_inputs(circ, standard_depolarizing_model())
    monkeypatch.setenv("QEC_B200_ENUM", "linear")
    gl = capi.Circuit(circ.nq, circ.nc, circ.ops)
    el = make_enumerator(capi, gl, order, idx, paulis)
    monkeypatch.setenv("QEC_B200_ENUM", "frame")
    gf = capi.Circuit(circ.nq, circ.nc, circ.ops)
    ef = make_enumerator(capi, gf, order, idx, paulis)
    n = el.num_paths()
    assert n == ef.num_paths()
    rng = np.random.default_rng(3)
    ranges = [(0, min(n, 70_000)), (max(0, n - 33_333), n), (0, 1), (n - 1, n)]
    ranges += [(int(lo), int(min(n, lo + rng.integers(1, 5000)))) for lo in rng.integers(0, n, size=6)]
    for lo, hi in ranges:
        bl, tl = el.range(lo, hi, want_tally=True)
        assert "enumerate:linear" in capi.last_kernel(), capi.last_kernel()
        _, tl2 = el.range(lo, hi, want_bits=False, want_tally=True)
        bl3, _ = el.range(lo, hi)
        bf, tf = ef.range(lo, hi, want_tally=True)
        assert "enumerate:linear" not in capi.last_kernel()
        assert np.array_equal(bl, bf), f"bits differ on [{lo}, {hi})"
        assert np.array_equal(bl3, bf)
        assert np.array_equal(tl, tf), f"tallies differ on [{lo}, {hi})"
        assert np.array_equal(tl2, tf)


def test_linear_kernels_on_random_circuits_with_clbits_out_of_time_order(capi, port, monkeypatch):
    """Random gate streams measure qubits repeatedly into arbitrary clbits (overwrites, clbits not
    in time order, never-measured clbits): the linear kernels must fall back to raw signatures
    where their read-out schedules need ordered chains, and still agree with explicit propagation
    (enumerator: frame kernels; sampler: port + oracle replay inside the kernels' own checks)."""
    from helpers import random_circuit
    rng = np.random.default_rng(11)
    model = standard_depolarizing_model(0.02)
    for trial in range(4):
        circ = random_circuit(rng, nq=6, nc=9, n_ops=70)
        idx, labels, paulis = noise_inputs(circ, model)
        if len(idx) < 3:
            continue
        for order in (1, 2):
            monkeypatch.setenv("QEC_B200_ENUM", "linear")
            gl = capi.Circuit(circ.nq, circ.nc, circ.ops)
            el = make_enumerator(capi, gl, order, idx, paulis)
            monkeypatch.setenv("QEC_B200_ENUM", "frame")
            gf = capi.Circuit(circ.nq, circ.nc, circ.ops)
            ef = make_enumerator(capi, gf, order, idx, paulis)
            n = min(el.num_paths(), 50_000)
            bl, tl = el.range(0, n, want_tally=True)
            assert "enumerate:linear" in capi.last_kernel()
            bf, tf = ef.range(0, n, want_tally=True)
            assert np.array_equal(bl, bf) and np.array_equal(tl, tf), f"trial {trial} order {order}"
        monkeypatch.delenv("QEC_B200_ENUM")
        _, _, lw, lp = sampler_inputs(circ, model)
        for mode in ("lanes16", "lanes", "warp", "cta"):
            monkeypatch.setenv("QEC_B200_LIN_MODE", mode)
            g = capi.Circuit(circ.nq, circ.nc, circ.ops)
            sm, _ = make_sampler(capi, g, idx, labels, lw, lp, seed=trial)
            n_shots = 6_000
            bits, tally, cnt, ev = sm.sample(0, n_shots, want_tally=True, fault_cap=96)
            assert "sample:linear" in capi.last_kernel()
            tab = sm.tables()
            paths = [([], []) for _ in range(n_shots)]
            for w in range((n_shots + 31) // 32):
                want_ev = port.sample_word(tab, w, min(32, n_shots - 32 * w))
                assert np.array_equal(ev[w, :cnt[w]], want_ev)
                for e in want_ev:
                    loc, bit, dig = int(e) >> 14, (int(e) >> 9) & 31, int(e) & 511
                    paths[32 * w + bit][0].append(idx[loc])
                    paths[32 * w + bit][1].append(paulis[loc][dig])
            want = port.propagate_batch(circ.nq, circ.nc, circ.ops, paths)
            assert np.array_equal(capi.unpack_bits(bits, n_shots), want), f"trial {trial} mode {mode}"
            assert np.array_equal(tally[:circ.nc], want.sum(axis=0).astype(np.uint64))
        monkeypatch.delenv("QEC_B200_LIN_MODE")


@pytest.mark.parametrize("flush", [1, 3])
def test_register_tally_counters_flush_mid_launch(capi, monkeypatch, flush):
    """The lane kernels keep the per-clbit counts of whole batches in registers and add them to
    the tally row every N batches (N = 448 by default: more batches per warp than any test
    launch has).  With N = 1 and 3 (QEC_B200_LANES_FLUSH) the mid-launch flush runs many times per
    warp: tallies must equal those of the default and the column sums of the bits, for the
    sampler (aligned and unaligned windows) and the enumerator (rows not 16-byte aligned)."""
    circ = _circuit("surf5")
    model = standard_depolarizing_model(2e-3)
    idx, labels, lw, lp = sampler_inputs(circ, model)
    _, _, paulis = noise_inputs(circ, model)
    n = 148 * 9 * 1024 * 5 + 32 * 7          # several batches per warp, a ragged end
    n_en, lo_en = 2_000_000 + 17, 1_000_003   # n_words of the window is not a multiple of 4

    def run():
        g = capi.Circuit(circ.nq, circ.nc, circ.ops)
        sm, _ = make_sampler(capi, g, idx, labels, lw, lp, seed=11)
        bits, tally, _, _ = sm.sample(0, n, want_tally=True)
        assert "lane-streams" in capi.last_kernel()
        bits_u, tally_u, _, _ = sm.sample(4096 + 32, n, want_tally=True)   # rows off the 16-byte grid
        en = make_enumerator(capi, g, 2, idx, paulis)
        be, te = en.range(lo_en, lo_en + n_en, want_tally=True)
        assert "enumerate:linear" in capi.last_kernel()
        return bits, tally, bits_u, tally_u, be, te

    ref_out = run()
    monkeypatch.setenv("QEC_B200_LANES_FLUSH", str(flush))
    out = run()
    for a, b in zip(ref_out, out):
        assert np.array_equal(a, b)
    nc = circ.nc
    for bits, tally, count in ((out[0], out[1], n), (out[2], out[3], n), (out[4], out[5], n_en)):
        ones = np.array([int(np.unpackbits(bits[c].view(np.uint8)).sum()) for c in range(nc)], dtype=np.uint64)
        assert np.array_equal(tally[:nc], ones)
        assert int(tally[nc + 1]) == count


def test_lane_sampler_with_a_long_chain_program(capi, port):
    """Repetition code, 100 rounds: 203 clbits in two chains of 100 rows, so one lane group's
    read-out queue (100 passes) is longer than the unrolled trips of the fast path hold."""
    _check_sampler_against_port_and_oracle(capi, port, "rep100", 0.004, 6_000, 2048 + 64)
    assert "lane-streams" in capi.last_kernel()
