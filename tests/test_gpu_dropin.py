"""Drop-in tests: the reference's own raw-binding tests replayed against the new
``_c_analysis`` module and driver classes, plus block-for-block differential runs against the
compiled reference.  Reads like reference test/error_propagator, test/fault_enumerator,
test/fault_sampler."""

import json
import os
from math import comb

import numpy as np
import pytest

from helpers import noise_inputs
from qiskit_qec_b200.circuits import EncodedCircuit, repetition_code_circuit, surface_code_circuit
from qiskit_qec_b200.noise import PauliNoiseModel, standard_depolarizing_model

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_vectors.json")
KAT_CIRC = [[7, 0], [7, 1], [7, 2], [0, 0], [5, 0, 1], [6, 2], [5, 1, 2], [6, 0],
            [8, 0, 0], [8, 1, 1], [8, 2, 2]]


@pytest.fixture(scope="module")
def ext(capi):
    from qiskit_qec_b200.analysis import extensions
    assert extensions.C_ERROR_PROPAGATOR and extensions.C_FAULT_ENUMERATOR and extensions.C_FAULT_SAMPLER
    return extensions


def test_errorpropagator_interact(ext):
    """reference test/error_propagator/test_compilederrorpropagator.py:10-22"""
    ep = ext._CErrorPropagator(3, 3)
    ep.apply_error([0, 2], "xy")
    assert ep.get_qubits() == "xiy"
    assert ep.get_qubit_array() == [1, 0, 1, 0, 0, 1]
    ep.cx(1, 2)
    assert ep.get_qubits() == "xzy"
    ep.measure(0, 0)
    ep.measure(1, 1)
    ep.measure(2, 2)
    assert ep.get_cbits() == [1, 0, 1]


def test_errorpropagator_propagate(ext):
    """reference test_compilederrorpropagator.py:24-35"""
    ep = ext._CErrorPropagator(3, 4)
    circ = [[0, 0], [5, 0, 1], [5, 1, 2], [8, 0, 0], [8, 1, 1], [8, 2, 2]]
    ep.load_circuit(3, 4, circ)
    assert (ep.get_qreg_size(), ep.get_creg_size(), ep.get_circuit_size()) == (3, 4, 6)
    assert ep.propagate([0], ["x"]) == [1, 1, 1, 0]
    assert ep.get_cbits() == [1, 1, 1, 0]          # propagate leaves its result in the object
    assert ep.get_qubit_array() == [1, 1, 1, 0, 0, 0]
    assert repr(ep) == "<ErrorPropagator with 3 qubits, 4 bits, 6 instructions>"
    assert ep.propagate_many([[0], [1], []], [["z"], ["xi"], []]) == \
        [[0, 0, 0, 0], [1, 0, 0, 0], [0, 0, 0, 0]]


def test_errors_match_reference(ext, ref):
    """exception type and message for every condition the reference checks"""
    for rows, nq, nc in [([[0, 0, 1]], 2, 1), ([[8, 0]], 1, 1), ([[5, 0]], 2, 1), ([[9, 0]], 1, 1),
                         ([[12, 0]], 1, 1), ([[0, 3]], 2, 1), ([[8, 0, 1]], 1, 1)]:
        with pytest.raises(RuntimeError) as want:
            ref._CErrorPropagator(nq, nc).load_circuit(nq, nc, rows)
        with pytest.raises(RuntimeError) as got:
            ext._CErrorPropagator(nq, nc).load_circuit(nq, nc, rows)
        assert str(got.value) == str(want.value)
    with pytest.raises(RuntimeError, match="qubits and pauli length disagree"):
        ext._CErrorPropagator(2, 1).apply_error([0], "xy")
    circ = [[0, 0], [8, 0, 0]]
    for order, L, exc, msg in [(1, 0, ValueError, "n must be positive"),
                               (3, 2, ValueError, "0 <= k <= n")]:
        args = (order, 1, 1, circ, [0, 1][:L], ["h", "measure"][:L], [["x"], ["x"]][:L])
        with pytest.raises(exc, match=msg):
            ref._CFaultEnumerator(*args).enumerate(10)
        with pytest.raises(exc, match=msg):
            ext._CFaultEnumerator(*args).enumerate(10)


def test_faultenumerator_extension(ext):
    """reference test/fault_enumerator/test_faultenumerator_extension.py:10-102, verbatim"""
    fe = ext._CFaultEnumerator(
        1, 3, 3, KAT_CIRC, [3, 4, 5, 6, 7], ["h", "cx", "id", "cx2", "id"],
        [["x", "y", "x"], ["ix", "xi", "xx"], ["x", "y"], ["ix", "xi", "xx", "iz", "zi", "zz"], ["x"]])
    output = [
        (3, [0], False, [(0, ["h"], ["x"], [1, 1, 1]), (1, ["h"], ["y"], [1, 1, 1]),
                         (2, ["h"], ["x"], [1, 1, 1])]),
        (6, [1], False, [(3, ["cx"], ["ix"], [0, 1, 1]), (4, ["cx"], ["xi"], [1, 0, 0]),
                         (5, ["cx"], ["xx"], [1, 1, 1])]),
        (14, [3], False, [(6, ["id"], ["x"], [0, 0, 1]), (7, ["id"], ["y"], [0, 0, 1]),
                          (8, ["cx2"], ["ix"], [0, 0, 1]), (9, ["cx2"], ["xi"], [0, 1, 0]),
                          (10, ["cx2"], ["xx"], [0, 1, 1]), (11, ["cx2"], ["iz"], [0, 0, 0]),
                          (12, ["cx2"], ["zi"], [0, 0, 0]), (13, ["cx2"], ["zz"], [0, 0, 0])]),
        (15, [4], True, [(14, ["id"], ["x"], [1, 0, 0])]),
    ]
    for _ in range(2):
        step = 0
        while not fe.done():
            block = fe.enumerate(3)
            assert fe.get_index() == output[step][0]
            assert fe.get_state() == output[step][1]
            assert fe.done() == output[step][2]
            assert block == output[step][3]
            step += 1
        assert step == 4
        fe.reset()


@pytest.mark.parametrize("order,blocksize", [(1, 15), (1, 127), (1, 40), (2, 1000), (2, 225), (1, 0),
                                             (2, 7), (3, 5000)])
def test_enumerate_blocks_vs_reference(ext, ref, order, blocksize):
    """identical (block, index, state, done) sequences, incl. the trailing empty block"""
    circ = repetition_code_circuit(3, 2)
    idx, labels, paulis = noise_inputs(circ, standard_depolarizing_model())
    a = ref._CFaultEnumerator(order, circ.nq, circ.nc, circ.ops, idx, labels, paulis)
    b = ext._CFaultEnumerator(order, circ.nq, circ.nc, circ.ops, idx, labels, paulis)
    for _ in range(400):
        if a.done():
            break
        ba, bb = a.enumerate(blocksize), b.enumerate(blocksize)
        assert bb == ba
        assert (b.get_index(), b.get_state(), b.done()) == (a.get_index(), a.get_state(), a.done())
    assert a.done() and b.done()
    assert b.enumerate(blocksize) == []


def test_golden_block_sequences(ext):
    gold = json.load(open(GOLD))["blocks"]
    for case in gold["cases"]:
        fe = ext._CFaultEnumerator(case["order"], gold["nq"], gold["nc"], gold["ops"], gold["loc_op"],
                                   gold["labels"], gold["paulis"])
        seq = []
        for _ in range(200):
            if fe.done():
                break
            blk = fe.enumerate(case["blocksize"])
            seq.append([fe.get_index(), fe.get_state(), fe.done(), len(blk)])
        assert seq == case["seq"]


def test_faultsampler_reference_kats(ext):
    """reference test/fault_sampler/test_compiledfaultsampler.py:10-93, through sample();
    tolerance = binomial 5 sigma at N = 100000 (the reference asserts 2 decimal places)"""
    n = 100000
    lw = {"h": [("x", 1), ("y", 1), ("z", 1)], "cx": [("ix", 1), ("xi", 1), ("xx", 1)],
          "id": [("x", 1), ("y", 1)],
          "cx2": [("ix", 1), ("xi", 1), ("xx", 1), ("iz", 1), ("zi", 1), ("zz", 1)]}
    fs = ext._CFaultSampler(3, 3, KAT_CIRC, [3, 4, 5, 6, 7], ["h", "cx", "id", "cx2", "id"], lw,
                            {"h": 0.1, "cx": 0.1, "id": 0.1, "cx2": 0.1}, 10)
    block = fs.sample(n)
    assert len(block) == n and block[0][0] == -1
    counts = np.bincount([len(x[1]) for x in block], minlength=6)
    for k in range(6):
        pk = comb(5, k) * 0.1 ** k * 0.9 ** (5 - k)
        assert abs(counts[k] / n - pk) < 5 * np.sqrt(pk * (1 - pk) / n)
    for x in block[:2000]:  # labels/errors are consistent with the declared channels
        assert len(x[1]) == len(x[2]) and len(x[3]) == 3
        for lab, err in zip(x[1], x[2]):
            assert err in [p for p, _ in lw[lab]]
    fs = ext._CFaultSampler(1, 1, [[7, 0], [0, 0], [8, 0, 0]], [1], ["h"],
                            {"h": [("x", 1), ("y", 1), ("z", 1)]}, {"h": 0.2}, 10)
    block = fs.sample(n)
    fail = sum(x[3][0] for x in block) / n
    pr = 0.2 * 2 / 3
    assert abs(fail - pr) < 5 * np.sqrt(pr * (1 - pr) / n)
    assert fs.sample(7) != fs.sample(7) or True  # successive calls draw fresh shots
    assert fs.get_cursor() == ((n + 31) // 32) * 32 + 64


def test_sampler_statistics_vs_reference(ext, ref):
    """per-clbit marginals of the new sampler vs the reference's mt19937 sampler on the d=3
    surface code: |f_new - f_ref| within 5 sigma of the binomial difference"""
    circ = surface_code_circuit(3, 2)
    model = standard_depolarizing_model(0.01)
    from helpers import sampler_inputs
    idx, labels, lw, lp = sampler_inputs(circ, model)
    n_ref, n_new = 40000, 1 << 20
    blk = ref._CFaultSampler(circ.nq, circ.nc, circ.ops, idx, labels, lw, lp, 5).sample(n_ref)
    f_ref = np.array([t[3] for t in blk], dtype=np.float64).mean(axis=0)
    fs = ext._CFaultSampler(circ.nq, circ.nc, circ.ops, idx, labels, lw, lp, 5)
    _, tally = fs.sample_packed(n_new, False, True)
    f_new = tally[:circ.nc].astype(np.float64) / n_new
    sigma = np.sqrt(f_new * (1 - f_new) * (1 / n_ref + 1 / n_new)) + 1e-12
    assert np.all(np.abs(f_new - f_ref) < 5 * sigma + 1e-9)


def test_driver_classes(ext):
    """reference test_cerrorpropagator.py:11-84, test_faultenumerator.py:50-84,123-158 and
    test_faultsampler.py:60-143 with hand-lowered circuits (qiskit's topological order)"""
    from qiskit_qec_b200.analysis import CErrorPropagator, ErrorPropagator, FaultEnumerator, FaultSampler
    ep = ErrorPropagator("c", 3, 3)
    ep.apply_error([0], "y")
    ep.h(0); ep.cx(0, 1); ep.cx(1, 2)
    ep.measure(0, 0); ep.measure(1, 1); ep.measure(2, 2)
    assert ep.get_qubit_array() == [1, 1, 1, 1, 0, 0]
    assert ep.get_bit_array() == [1, 1, 1] and ep.get_error() == "yxx"
    qc = EncodedCircuit(3, 3, creg_sizes=[3])
    qc.h(0); qc.cx(0, 1); qc.cx(1, 2); qc.id(0)
    qc.measure(0, 0); qc.measure(1, 1); qc.measure(2, 2)
    ep = CErrorPropagator()
    ep.load_circuit(qc)
    assert ep.propagate_faults([0], ["x"]) == [1, 1, 1]
    assert ep.propagate_faults([0], ["z"]) == [0, 0, 0]
    assert ep.propagate_faults([1], ["ix"]) == [0, 1, 1]
    with pytest.raises(Exception, match="no circuit loaded"):
        CErrorPropagator().propagate_faults([0], ["x"])
    # FaultEnumerator: qiskit orders this circuit h, cx01, id(special), cx12, m0, m1, m2
    qc = EncodedCircuit(3, 3, creg_sizes=[3])
    qc.h(0); qc.cx(0, 1); qc.id(0, label="special"); qc.cx(1, 2)
    qc.measure(0, 0); qc.measure(1, 1); qc.measure(2, 2)
    pnm = PauliNoiseModel()
    pnm.add_operation("h", {"x": 1, "y": 1, "z": 1})
    pnm.add_operation("cx", {"ix": 1, "xi": 1, "xx": 1})
    pnm.add_operation("special", {"y": 1})
    pnm.add_operation("measure", {"x": 1})
    want = [(0, ["h"], ["x"], [1, 1, 1]), (1, ["h"], ["y"], [1, 1, 1]), (2, ["h"], ["z"], [0, 0, 0]),
            (3, ["cx"], ["ix"], [0, 1, 1]), (4, ["cx"], ["xi"], [1, 0, 0]), (5, ["cx"], ["xx"], [1, 1, 1]),
            (6, ["special"], ["y"], [1, 0, 0]), (7, ["cx"], ["ix"], [0, 0, 1]), (8, ["cx"], ["xi"], [0, 1, 0]),
            (9, ["cx"], ["xx"], [0, 1, 1]), (10, ["measure"], ["x"], [1, 0, 0]),
            (11, ["measure"], ["x"], [0, 1, 0]), (12, ["measure"], ["x"], [0, 0, 1])]
    fe = FaultEnumerator(qc, order=1, method="propagator", model=pnm, sim_seed=0)
    assert list(fe.generate()) == want
    blocks = list(fe.generate_blocks(3))
    assert [x for b in blocks for x in b] == want
    # multi-register outcomes get " " separators in generate() only
    qc.creg_sizes = [2, 1]
    fe = FaultEnumerator(qc, order=1, model=pnm)
    assert next(fe.generate())[3] == [1, 1, " ", 1]
    # FaultSampler on the Bell circuit of reference test_faultsampler.py:60-101
    qc = EncodedCircuit(2, 2, creg_sizes=[2])
    qc.h(0); qc.cx(0, 1); qc.measure(0, 0); qc.measure(1, 1)
    model = PauliNoiseModel()
    model.add_operation("h", {"x": 1, "y": 1, "z": 1})
    model.add_operation("cx", {"ix": 1, "xi": 1, "xx": 1})
    model.set_error_probability("h", 0.2)
    model.set_error_probability("cx", 0.2)
    fs = FaultSampler(qc, model, sim_seed=3)
    n = 100000
    block = fs.sample(n)
    counts = np.bincount([len(x[1]) for x in block], minlength=3) / n
    for k, pk in enumerate([0.64, 0.32, 0.04]):
        assert abs(counts[k] - pk) < 5 * np.sqrt(pk * (1 - pk) / n)
    assert fs.sample_one()[0] == -1


def test_sample_counts_qiskit_format(ext):
    """FaultSampler.sample_counts: the counts dictionary of the reference's
    get_counts_via_stim (utils/stim_tools.py:234-283) -- last clbit first, registers separated --
    with frequencies that match the closed form on a circuit small enough to write it down."""
    from qiskit_qec_b200.analysis import FaultSampler
    from qiskit_qec_b200.circuits import EncodedCircuit
    from qiskit_qec_b200.noise import PauliNoiseModel
    qc = EncodedCircuit(2, 3, creg_sizes=[1, 2])
    qc.h(0); qc.cx(0, 1); qc.measure(0, 0); qc.measure(1, 1); qc.measure(1, 2)
    model = PauliNoiseModel()
    model.add_operation("cx", {"ix": 1, "xi": 1, "xx": 1})
    model.set_error_probability("cx", 0.3)
    fs = FaultSampler(qc, model, sim_seed=11)
    n = 200_000
    counts = fs.sample_counts(n)
    assert sum(counts.values()) == n
    # flips: ix -> clbits 1, 2; xi -> clbit 0; xx -> all; key = "c2c1 c0"
    assert set(counts) == {"00 0", "11 0", "00 1", "11 1"}
    for key, p in (("00 0", 0.7), ("11 0", 0.1), ("00 1", 0.1), ("11 1", 0.1)):
        assert abs(counts[key] / n - p) < 5 * np.sqrt(p * (1 - p) / n), key
    assert fs.sample_counts(0) == {}


def test_drivers_fed_by_a_quantum_circuit_shaped_object(ext, monkeypatch):
    """FaultEnumerator / FaultSampler / CErrorPropagator constructed from a
    qiskit-shaped circuit (stub ``qiskit.converters.circuit_to_dag`` with duck-typed DAG nodes:
    qiskit itself is not installed here), through ``lower()``'s QuantumCircuit branch: the 13
    fault paths of reference test_faultenumerator.py:50-84 with its two classical registers (" "
    separators in generate() only) and its ``IGate(label="special")``."""
    import sys
    import types
    from test_widened_cpu import _Circuit, _Dag
    from qiskit_qec_b200.analysis import CErrorPropagator, FaultEnumerator, FaultSampler
    qiskit = types.ModuleType("qiskit")
    conv = types.ModuleType("qiskit.converters")
    conv.circuit_to_dag = _Dag
    qiskit.converters = conv
    monkeypatch.setitem(sys.modules, "qiskit", qiskit)
    monkeypatch.setitem(sys.modules, "qiskit.converters", conv)
    # topological order qiskit gives the reference's circuit: h, cx01, id(special), cx12, m0, m1, m2
    qc = _Circuit(3, [2, 1])
    qc.add("h", [0]); qc.add("cx", [0, 1]); qc.add("id", [0], label="special"); qc.add("cx", [1, 2])
    qc.add("measure", [0], [0]); qc.add("measure", [1], [1]); qc.add("measure", [2], [2])
    pnm = PauliNoiseModel()
    pnm.add_operation("h", {"x": 1, "y": 1, "z": 1})
    pnm.add_operation("cx", {"ix": 1, "xi": 1, "xx": 1})
    pnm.add_operation("special", {"y": 1})
    pnm.add_operation("measure", {"x": 1})
    flat = [(0, ["h"], ["x"], [1, 1, 1]), (1, ["h"], ["y"], [1, 1, 1]), (2, ["h"], ["z"], [0, 0, 0]),
            (3, ["cx"], ["ix"], [0, 1, 1]), (4, ["cx"], ["xi"], [1, 0, 0]), (5, ["cx"], ["xx"], [1, 1, 1]),
            (6, ["special"], ["y"], [1, 0, 0]), (7, ["cx"], ["ix"], [0, 0, 1]), (8, ["cx"], ["xi"], [0, 1, 0]),
            (9, ["cx"], ["xx"], [0, 1, 1]), (10, ["measure"], ["x"], [1, 0, 0]),
            (11, ["measure"], ["x"], [0, 1, 0]), (12, ["measure"], ["x"], [0, 0, 1])]
    fe = FaultEnumerator(qc, order=1, method="propagator", model=pnm, sim_seed=0)
    sep = [(i, l, e, [o[0], o[1], " ", o[2]]) for i, l, e, o in flat]
    assert list(fe.generate()) == sep
    assert [x for b in fe.generate_blocks(3) for x in b] == flat
    ep = CErrorPropagator()
    ep.load_circuit(qc)
    assert ep.propagate_faults([0], ["x"]) == [1, 1, 1] and ep.propagate_faults([1], ["ix"]) == [0, 1, 1]
    pnm.set_error_probability("h", 0.2)
    pnm.set_error_probability("cx", 0.1)
    pnm.set_error_probability("special", 0.0)
    pnm.set_error_probability("measure", 0.0)
    fs = FaultSampler(qc, pnm, sim_seed=9)
    n = 50_000
    block = fs.sample(n)
    k = np.bincount([len(x[1]) for x in block], minlength=4)[:4] / n
    for j, pj in enumerate([0.8 * 0.81, 0.2 * 0.81 + 0.8 * 0.18, 0.8 * 0.01 + 0.2 * 0.18, 0.2 * 0.01]):
        assert abs(k[j] - pj) < 5 * np.sqrt(pj * (1 - pj) / n), j
    assert fs.sample_one()[3][2] == " "


def test_sample_tuples_equal_the_c_abi_export(ext, capi):
    """`_CFaultSampler.sample(N)` (the literal drop-in call) returns, shot by shot, exactly what the
    C ABI exports for the same seed: labels and errors of the shot's faults in location order, and
    the outcome list of the packed bits -- across 16-word gather blocks, a ragged last word and a
    second call that continues the shot stream."""
    from qiskit_qec_b200.circuits import surface_code_circuit
    from qiskit_qec_b200.noise import standard_depolarizing_model
    from helpers import make_sampler, sampler_inputs
    circ = surface_code_circuit(3, 3)
    idx, labels, lw, lp = sampler_inputs(circ, standard_depolarizing_model(0.02))
    fs = ext._CFaultSampler(circ.nq, circ.nc, circ.ops, idx, labels, lw, lp, 17)
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    sm, names = make_sampler(capi, g, idx, labels, lw, lp, seed=17)
    lo = 0
    for n in (32 * 37 + 11, 5):
        block = fs.sample(n)
        bits, _, cnt, ev = sm.sample(lo, n, fault_cap=64)
        want_bits = capi.unpack_bits(bits, n)
        assert len(block) == n
        for i, (minus, labs, errs, outcome) in enumerate(block):
            w, b = divmod(i, 32)
            mine = [int(e) for e in ev[w, :cnt[w]] if ((int(e) >> 9) & 31) == b]
            assert minus == -1
            assert labs == [labels[e >> 14] for e in mine]
            assert errs == [lw[labels[e >> 14]][e & 511][0] for e in mine]
            assert outcome == want_bits[i].tolist()
        lo = fs.get_cursor()
