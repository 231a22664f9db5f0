"""BASELINE.json's full sizes, checked through size-independent properties.

The frame map is GF(2)-linear in the injected faults (reference errorpropagator.cpp:211-313:
every gate is linear, faults are XORed in), so the order-1 signatures -- taken here from the
UNMODIFIED reference enumerator -- predict exactly
  * the per-clbit marginal counts of the complete order-2 enumeration (cfg3, 21 740 940 paths),
  * the per-clbit flip probabilities of the sampler (cfg2, 1e8 shots; binomial 5 sigma).
Plus shard invariance: the tallies of one launch equal the sum over four shards, bit for bit,
and the packed outcome bits agree with the tallies.
"""
import numpy as np
import pytest
import torch

from helpers import make_enumerator, make_sampler, noise_inputs, sampler_inputs
from qiskit_qec_b200.circuits import surface_code_circuit
from qiskit_qec_b200.noise import standard_depolarizing_model
from qiskit_qec_b200.sharding import shard_range

pytestmark = pytest.mark.gpu
P = 1e-3


@pytest.fixture(scope="module")
def order1(ref):
    """a[l][c] = number of Paulis of location l whose single fault flips clbit c; n[l]"""
    circ = surface_code_circuit(5, 5)
    idx, labels, paulis = noise_inputs(circ, standard_depolarizing_model(P))
    fe = ref._CFaultEnumerator(1, circ.nq, circ.nc, circ.ops, idx, labels, paulis)
    block = fe.enumerate(10 ** 6)
    assert len(block) == sum(len(p) for p in paulis) == 6601
    a = np.zeros((len(idx), circ.nc), dtype=np.int64)
    k = 0
    for l, pl in enumerate(paulis):
        for _ in pl:
            a[l] += np.array(block[k][3], dtype=np.int64)
            k += 1
    n = np.array([len(pl) for pl in paulis], dtype=np.int64)
    return circ, idx, labels, paulis, a, n


def test_cfg3_order2_marginals_equal_linear_prediction(capi, order1):
    circ, idx, labels, paulis, a, n = order1
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    en = make_enumerator(capi, g, 2, idx, paulis)
    total = en.num_paths()
    assert total == 21_740_940
    dev = torch.device("cuda", 0)
    tally = torch.zeros(capi.tally_len(circ.nc), dtype=torch.int64, device=dev)
    bits = torch.empty((circ.nc, (total + 31) // 32), dtype=torch.int32, device=dev)
    en.range_dev(0, total, bits.data_ptr(), tally.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    t = tally.cpu().numpy()
    # pairs l1 < l2: bit c of the XOR is set iff exactly one of the two Paulis flips it
    A, N = a.sum(axis=0), n.sum()
    want = (A * N - (a * n[:, None]).sum(axis=0)) - (A * A - (a * a).sum(axis=0))
    np.testing.assert_array_equal(t[:circ.nc], want)
    assert t[circ.nc + 1] == total
    # the packed bits carry the same counts (popcount per clbit row)
    rows = bits.view(torch.uint8)
    lut = torch.tensor([bin(i).count("1") for i in range(256)], dtype=torch.int64, device=dev)
    pop = lut[rows.long()].sum(dim=1).cpu().numpy()
    np.testing.assert_array_equal(pop, want)
    # sharded over 4 ranks' ranges: identical tallies
    acc = torch.zeros_like(tally)
    for r in range(4):
        lo, hi = shard_range(total, r, 4)
        en.range_dev(lo, hi, 0, acc.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    np.testing.assert_array_equal(acc.cpu().numpy(), t)


def test_cfg2_sampler_1e8_shots_statistics_and_shard_invariance(capi, order1):
    circ, idx, labels, paulis, a, n = order1
    _, _, lw, lp = sampler_inputs(circ, standard_depolarizing_model(P))
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    sm, _ = make_sampler(capi, g, idx, labels, lw, lp, seed=2026)
    n_shots = 100_000_000
    dev = torch.device("cuda", 0)
    stream = torch.cuda.current_stream().cuda_stream
    tally = torch.zeros(capi.tally_len(circ.nc), dtype=torch.int64, device=dev)
    sm.sample_dev(0, n_shots, 0, tally.data_ptr(), stream)
    torch.cuda.synchronize()
    sm.check()
    t = tally.cpu().numpy()
    nc = circ.nc
    assert t[nc + 1] == n_shots and t[nc + 2:].sum() == n_shots
    # faults per shot: Binomial(761, p)
    mean = (t[nc + 2:] * np.arange(64)).sum() / n_shots
    mu = len(idx) * P
    assert abs(mean - mu) < 5 * np.sqrt(mu * (1 - P) / n_shots)
    p0 = t[nc + 2] / n_shots
    q0 = (1 - P) ** len(idx)
    assert abs(p0 - q0) < 5 * np.sqrt(q0 * (1 - q0) / n_shots)
    # clbit c flips iff an odd number of independent location faults flip it
    q = np.prod(1.0 - 2.0 * P * a / n[:, None], axis=0)
    pc = (1.0 - q) / 2.0
    sig = np.sqrt(pc * (1 - pc) / n_shots)
    z = (t[:nc] / n_shots - pc) / sig
    assert np.abs(z).max() < 5.0, f"worst clbit z = {z[np.abs(z).argmax()]:.2f}"
    # four shards == one launch, bit for bit (counter-based RNG keyed by the global shot id)
    acc = torch.zeros_like(tally)
    for r in range(4):
        lo, hi = shard_range(n_shots, r, 4)
        sm.sample_dev(lo, hi - lo, 0, acc.data_ptr(), stream)
    torch.cuda.synchronize()
    sm.check()
    np.testing.assert_array_equal(acc.cpu().numpy(), t)
    # packed bits of a 2e7-shot window agree with its tallies
    m = 20_000_000
    tw = torch.zeros_like(tally)
    bits = torch.empty((nc, (m + 31) // 32), dtype=torch.int32, device=dev)
    sm.sample_dev(32 * 1000, m, bits.data_ptr(), tw.data_ptr(), stream)
    torch.cuda.synchronize()
    lut = torch.tensor([bin(i).count("1") for i in range(256)], dtype=torch.int64, device=dev)
    pop = torch.stack([lut[bits[c].view(torch.uint8).long()].sum() for c in range(nc)]).cpu().numpy()
    np.testing.assert_array_equal(pop, tw.cpu().numpy()[:nc])


def test_host_entry_points_stream_in_chunks(capi, order1, monkeypatch):
    """The host-pointer entry points stream their outcomes back chunk by chunk on two streams
    (qec_capi.cu pipelined_out); with 1 MiB chunks a 1e6-shot call is ~31 chunks, the last one
    ragged.  Results must equal the single-launch device path bit for bit."""
    circ, idx, labels, paulis, a, n = order1
    _, _, lw, lp = sampler_inputs(circ, standard_depolarizing_model(P))
    g = capi.Circuit(circ.nq, circ.nc, circ.ops)
    sm, _ = make_sampler(capi, g, idx, labels, lw, lp, seed=5)
    dev = torch.device("cuda", 0)
    stream = torch.cuda.current_stream().cuda_stream
    n_shots, lo = 1_000_003, 64 * 12345
    n_words = (n_shots + 31) // 32
    tally = torch.zeros(capi.tally_len(circ.nc), dtype=torch.int64, device=dev)
    bits = torch.empty((circ.nc, n_words), dtype=torch.int32, device=dev)
    sm.sample_dev(lo, n_shots, bits.data_ptr(), tally.data_ptr(), stream)
    torch.cuda.synchronize()
    monkeypatch.setenv("QEC_B200_PIPE_MB", "1")
    hbits, htally, _, _ = sm.sample(lo, n_shots, want_tally=True)
    np.testing.assert_array_equal(hbits.view(np.int32).reshape(circ.nc, n_words), bits.cpu().numpy())
    np.testing.assert_array_equal(htally.astype(np.int64), tally.cpu().numpy())
    # enumerator: a ragged window of order-2 paths
    en = make_enumerator(capi, g, 2, idx, paulis)
    plo, phi = 1_000_001, 1_000_001 + 2_000_037
    pw = (phi - plo + 31) // 32
    etally = torch.zeros_like(tally)
    ebits = torch.empty((circ.nc, pw), dtype=torch.int32, device=dev)
    en.range_dev(plo, phi, ebits.data_ptr(), etally.data_ptr(), stream)
    torch.cuda.synchronize()
    hb, ht = en.range(plo, phi, want_tally=True)
    np.testing.assert_array_equal(hb.view(np.int32).reshape(circ.nc, pw), ebits.cpu().numpy())
    np.testing.assert_array_equal(ht.astype(np.int64), etally.cpu().numpy())
