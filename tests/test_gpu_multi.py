"""Multi-GPU handles of the product (qec_multi_*, one process): N devices give the bits and the
tallies of one device.  Runs with every device count from 1 to what the box has (the 1-device
case still goes through the sharding, host-thread, row-pitch and reduce code)."""

import numpy as np
import pytest

from helpers import make_enumerator, make_sampler, noise_inputs, sampler_inputs
from qiskit_qec_b200.circuits import surface_code_circuit
from qiskit_qec_b200.noise import standard_depolarizing_model

pytestmark = pytest.mark.gpu


def _device_sets(capi):
    n = capi.device_count()
    return [list(range(k)) for k in sorted({1, min(2, n), min(3, n), n})]


def _sampler_args(capi, circ, model):
    idx, labels, lw, lp = sampler_inputs(circ, model)
    names = list(dict.fromkeys(labels))
    cls = {n: i for i, n in enumerate(names)}
    return (idx, [cls[lab] for lab in labels], [lp.get(n, 0.0) for n in names],
            [[capi.pauli_code(p) for p, _ in lw[n]] for n in names], [[w for _, w in lw[n]] for n in names])


@pytest.mark.parametrize("d,T,p,n_shots,lo", [(5, 5, 2e-3, 3 * 4096 * 5 + 32 * 7 + 5, 64), (11, 3, 1e-3, 40_003, 0)])
def test_multi_sampler_equals_one_gpu(capi, d, T, p, n_shots, lo):
    circ = surface_code_circuit(d, T)
    model = standard_depolarizing_model(p)
    idx, labels, lw, lp = sampler_inputs(circ, model)
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    sm, _ = make_sampler(capi, g, idx, labels, lw, lp, seed=11)
    bits, tally, _, _ = sm.sample(lo, n_shots, want_tally=True)
    for devs in _device_sets(capi):
        ms = capi.MultiSampler(devs, circ.nq, circ.nc, circ.ops, *_sampler_args(capi, circ, model), seed=11)
        b, t = ms.sample(lo, n_shots, True, True)
        assert np.array_equal(b, bits) and np.array_equal(t, tally), devs
        _, t2 = ms.sample(lo, n_shots, False, True)
        t3, ms_dev = ms.tally_dev(lo, n_shots)
        assert np.array_equal(t2, tally) and np.array_equal(t3, tally) and ms_dev > 0, devs
        ms.close()


@pytest.mark.parametrize("order,lo,hi", [(1, 0, None), (2, 12_345, 12_345 + 200_000 + 17)])
def test_multi_enumerator_equals_one_gpu(capi, order, lo, hi):
    circ = surface_code_circuit(5, 5)
    idx, labels, paulis = noise_inputs(circ, standard_depolarizing_model())
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    en = make_enumerator(capi, g, order, idx, paulis)
    hi = en.num_paths() if hi is None else hi
    bits, tally = en.range(lo, hi, want_tally=True)
    codes = [[capi.pauli_code(p) for p in pl] for pl in paulis]
    for devs in _device_sets(capi):
        me = capi.MultiEnumerator(devs, circ.nq, circ.nc, circ.ops, order, idx, codes)
        assert me.num_paths() == en.num_paths()
        b, t = me.range(lo, hi, True, True)
        assert np.array_equal(b, bits) and np.array_equal(t, tally), devs
        t2, _ = me.tally_dev(lo, hi)
        assert np.array_equal(t2, tally), devs
        me.close()
    with pytest.raises(ValueError, match="listed twice"):
        capi.MultiEnumerator([0, 0], circ.nq, circ.nc, circ.ops, order, idx, codes)


def test_driver_devices_argument(capi):
    """FaultSampler / FaultEnumerator(devices=[...]): bulk calls on several GPUs == one GPU."""
    from qiskit_qec_b200.analysis import FaultEnumerator, FaultSampler
    devs = list(range(capi.device_count()))
    if len(devs) < 2:
        pytest.skip("one GPU visible: the multi handles are covered above with devices=[0]")
    circ = surface_code_circuit(3, 3)
    model = standard_depolarizing_model(0.01)
    one, many = FaultSampler(circ, model, sim_seed=4), FaultSampler(circ, model, sim_seed=4, devices=devs)
    for n in (50_000, 33):
        b1, t1 = one.sample_packed(n, True, True)
        b2, t2 = many.sample_packed(n, True, True)
        assert np.array_equal(np.asarray(b1), b2) and np.array_equal(np.asarray(t1), t2)
    e1, e2 = FaultEnumerator(circ, order=2, model=model), FaultEnumerator(circ, order=2, model=model, devices=devs)
    b1, t1 = e1.enumerate_packed(100, 90_001, want_tally=True)
    b2, t2 = e2.enumerate_packed(100, 90_001, want_tally=True)
    assert np.array_equal(np.asarray(b1), b2) and np.array_equal(np.asarray(t1), t2)
