"""GPU parity of the circuit-specialised (run-time compiled) kernels against the table-driven
kernels: same circuit, same faults -> identical packed outcomes and tallies, bit for bit.
The table-driven kernels are themselves checked against the reference in test_gpu_parity.py,
and with the default QEC_B200_JIT=1 those tests already run the specialised kernels wherever a
circuit is eligible, so both kernel families are pinned to the reference."""

import os

import numpy as np
import pytest

from qiskit_qec_b200.circuits import (arc_circuit, heavy_hex_27_arc_layout, repetition_code_circuit,
                                      surface_code_circuit)
from qiskit_qec_b200.noise import standard_depolarizing_model

from helpers import make_enumerator, make_sampler, noise_inputs, sampler_inputs

pytestmark = pytest.mark.gpu


class jit_env:
    def __init__(self, value):
        self.value = value

    def __enter__(self):
        # (these tests are about the frame-propagation kernels: keep the linear-response
        # enumerator, the default for long circuits, out of the way)
        self.old = os.environ.get("QEC_B200_JIT")
        self.old_enum = os.environ.get("QEC_B200_ENUM")
        os.environ["QEC_B200_JIT"] = self.value
        os.environ["QEC_B200_ENUM"] = "frame"

    def __exit__(self, *a):
        for key, old in (("QEC_B200_JIT", self.old), ("QEC_B200_ENUM", self.old_enum)):
            if old is None:
                os.environ.pop(key, None)
            else:
                os.environ[key] = old


def _circuit(name):
    if name == "rep":
        return repetition_code_circuit(3, 2)
    if name == "surf3":
        return surface_code_circuit(3, 3)
    if name == "surf5":
        return surface_code_circuit(5, 5)
    if name == "surf3x":
        return surface_code_circuit(3, 2, basis="x")
    links, sched, color = heavy_hex_27_arc_layout()
    return arc_circuit(links, 3, "zx", sched, color)


@pytest.mark.parametrize("name,order,windows", [
    ("rep", 1, None), ("rep", 2, None), ("rep", 3, None), ("surf3", 1, None), ("surf3", 2, None),
    ("surf3x", 2, None), ("arc", 2, 6), ("arc", 3, 6), ("surf5", 1, None), ("surf5", 2, 10),
    ("surf3", 3, 6),
])
def test_enumerator_specialised_equals_table_driven(capi, name, order, windows):
    circ = _circuit(name)
    idx, labels, paulis = noise_inputs(circ, standard_depolarizing_model())
    with jit_env("require"):
        gj = capi.Circuit(circ.nq, circ.nc, circ.ops)
        ej = make_enumerator(capi, gj, order, idx, paulis)
    with jit_env("0"):
        gt = capi.Circuit(circ.nq, circ.nc, circ.ops)
        et = make_enumerator(capi, gt, order, idx, paulis)
    n = ej.num_paths()
    rng = np.random.default_rng(17)
    if windows is None:
        ranges = [(0, n)] + [(int(lo), int(min(n, lo + rng.integers(0, 300))))
                             for lo in rng.integers(0, n, size=5)]
    else:
        span = 40_000
        ranges = [(0, span), (n - span, n)] + [(int(lo), int(lo) + span - int(rng.integers(0, 70)))
                                               for lo in rng.integers(0, n - span, size=windows)]
    for lo, hi in ranges:
        with jit_env("require"):
            bj, tj = ej.range(lo, hi, want_tally=True)
            assert "specialised" in capi.last_kernel()
            _, tj2 = ej.range(lo, hi, want_bits=False, want_tally=True)
            bj3, _ = ej.range(lo, hi)
        with jit_env("0"):
            bt, tt = et.range(lo, hi, want_tally=True)
            assert "table-driven" in capi.last_kernel()
        np.testing.assert_array_equal(bj, bt)
        np.testing.assert_array_equal(bj3, bt)
        np.testing.assert_array_equal(tj, tt)
        np.testing.assert_array_equal(tj2, tt)


@pytest.mark.parametrize("name,p,n_shots", [("rep", 0.05, 40_000), ("surf3", 0.01, 100_000),
                                            ("surf5", 1e-3, 1_000_000), ("arc", 2e-3, 200_000),
                                            ("surf5", 1e-4, 300_000)])
def test_sampler_specialised_equals_table_driven(capi, name, p, n_shots, monkeypatch):
    monkeypatch.setenv("QEC_B200_SAMPLER", "frame")  # the frame-propagation samplers, not the linear one
    circ = _circuit(name)
    idx, labels, lw, lp = sampler_inputs(circ, standard_depolarizing_model(p))
    out = {}
    for mode in ("require", "0"):
        with jit_env(mode):
            g = capi.Circuit(circ.nq, circ.nc, circ.ops)
            sm, _ = make_sampler(capi, g, idx, labels, lw, lp, seed=1234)
            res = []
            for lo, n in ((0, n_shots), (64 * 977, n_shots - 13), (32, 1)):
                bits, tally, _, _ = sm.sample(lo, n, want_tally=True)
                _, tally2, _, _ = sm.sample(lo, n, want_bits=False, want_tally=True)
                bits3, _, _, _ = sm.sample(lo, n)
                np.testing.assert_array_equal(tally, tally2)
                np.testing.assert_array_equal(bits, bits3)
                res.append((bits, tally))
            assert ("specialised" if mode == "require" else "table-driven") in capi.last_kernel()
            out[mode] = res
    for (bj, tj), (bt, tt) in zip(out["require"], out["0"]):
        np.testing.assert_array_equal(bj, bt)
        np.testing.assert_array_equal(tj, tt)


def test_ineligible_circuit_falls_back_to_table_driven(capi):
    """d=11 (482 frame words) cannot live in registers: table-driven kernel, and `require` fails."""
    circ = surface_code_circuit(11, 2)
    idx, labels, paulis = noise_inputs(circ, standard_depolarizing_model())
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    en = make_enumerator(capi, g, 1, idx, paulis)
    en.range(0, 1000)
    assert "table-driven" in capi.last_kernel() and "register" in capi.last_kernel()
    with jit_env("require"):
        with pytest.raises(RuntimeError, match="QEC_B200_JIT=require"):
            en.range(0, 1000)


@pytest.mark.parametrize("name,order,max_seg", [("surf3", 2, 3), ("surf5", 2, 8), ("arc", 3, 2)])
def test_enumerator_deferred_words_second_pass(capi, name, order, max_seg, monkeypatch):
    """QEC_B200_JIT_MAXSEG: lane words that need more unranking segments than the limit are
    listed by the specialised kernel and run by the table-driven kernel behind it; outcomes and
    tallies must not change."""
    circ = _circuit(name)
    idx, labels, paulis = noise_inputs(circ, standard_depolarizing_model())
    with jit_env("require"):
        g = capi.Circuit(circ.nq, circ.nc, circ.ops)
        en = make_enumerator(capi, g, order, idx, paulis)
        n = en.num_paths()
        lo, hi = max(0, n // 3 - 17), min(n, n // 3 + 150_000)
        monkeypatch.delenv("QEC_B200_JIT_MAXSEG", raising=False)
        b0, t0 = en.range(lo, hi, want_tally=True)
        monkeypatch.setenv("QEC_B200_JIT_MAXSEG", str(max_seg))
        b1, t1 = en.range(lo, hi, want_tally=True)
        _, t2 = en.range(lo, hi, want_bits=False, want_tally=True)
    np.testing.assert_array_equal(b1, b0)
    np.testing.assert_array_equal(t1, t0)
    np.testing.assert_array_equal(t2, t0)


def test_sampler_queue_overflow_reruns_on_table_driven(capi, monkeypatch):
    """A lane word that draws more faults than its shared-memory queue holds raises the
    overflow flag; the host entry point re-runs the call on the table-driven kernel (the RNG
    is counter based, so the re-run draws the same faults).  QEC_B200_JIT_CAPW shortens the
    queue to force this."""
    monkeypatch.setenv("QEC_B200_SAMPLER", "frame")
    circ = _circuit("surf3")
    idx, labels, lw, lp = sampler_inputs(circ, standard_depolarizing_model(0.02))
    n = 200_000
    with jit_env("0"):
        g0 = capi.Circuit(circ.nq, circ.nc, circ.ops)
        s0, _ = make_sampler(capi, g0, idx, labels, lw, lp, seed=77)
        b0, t0, _, _ = s0.sample(0, n, want_tally=True)
    monkeypatch.setenv("QEC_B200_JIT_CAPW", "12")
    with jit_env("1"):
        g1 = capi.Circuit(circ.nq, circ.nc, circ.ops)
        s1, _ = make_sampler(capi, g1, idx, labels, lw, lp, seed=77)
        b1, t1, _, _ = s1.sample(0, n, want_tally=True)
        assert "table-driven" in capi.last_kernel() and "more events" in capi.last_kernel()
    np.testing.assert_array_equal(b1, b0)
    np.testing.assert_array_equal(t1, t0)
