"""Multi-rank host logic on CPU: world_size-2 gloo.  Each rank enumerates its shard of the
order-2 path range with the CPU oracle (standing in for the GPU kernel, which the GPU tests
cover), the tallies are all-reduced, and the result must equal the single-range tally; the
concatenated packed outcomes must equal the single-range packing."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from helpers import noise_inputs
from oracle import port
from qiskit_qec_b200.circuits import repetition_code_circuit
from qiskit_qec_b200.noise import standard_depolarizing_model
from qiskit_qec_b200.sharding import allreduce_tally, shard_range, shard_shots


def test_shard_range_partitions_and_aligns():
    for total in (0, 1, 31, 32, 33, 1000, 21_740_940):
        for world in (1, 2, 3, 4, 8):
            pieces = [shard_range(total, r, world) for r in range(world)]
            assert pieces[0][0] == 0 and pieces[-1][1] == total
            for (lo, hi), (lo2, _) in zip(pieces, pieces[1:]):
                assert hi == lo2 and lo <= hi
            for lo, hi in pieces[:-1]:
                assert lo % 32 == 0 and hi % 32 == 0
            sizes = [hi - lo for lo, hi in pieces]
            assert max(sizes) - min(sizes) <= 32 or total < 32 * world
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)
    assert shard_shots(64, 100, 1, 2) == (64 + 64, 36)
    with pytest.raises(ValueError):
        shard_shots(5, 100, 0, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port_no, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port_no)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    circ = repetition_code_circuit(3, 2)
    idx, labels, paulis = noise_inputs(circ, standard_depolarizing_model())
    total, _ = port.enumerate_range(circ.nq, circ.nc, circ.ops, 2, idx, paulis, 0, 0)
    lo, hi = shard_range(total, rank, world)
    _, out = port.enumerate_range(circ.nq, circ.nc, circ.ops, 2, idx, paulis, lo, hi)
    tally = torch.zeros(circ.nc + 2, dtype=torch.int64)
    tally[:circ.nc] = torch.from_numpy(out.sum(axis=0).astype(np.int64))
    tally[circ.nc] = int((out.any(axis=1)).sum())
    tally[circ.nc + 1] = hi - lo
    allreduce_tally(tally)
    np.save(os.path.join(out_dir, f"tally{rank}.npy"), tally.numpy())
    np.save(os.path.join(out_dir, f"out{rank}.npy"), out)
    dist.destroy_process_group()


def test_two_rank_enumeration_equals_single_range(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    circ = repetition_code_circuit(3, 2)
    idx, labels, paulis = noise_inputs(circ, standard_depolarizing_model())
    total, full = port.enumerate_range(circ.nq, circ.nc, circ.ops, 2, idx, paulis)
    want = np.zeros(circ.nc + 2, dtype=np.int64)
    want[:circ.nc] = full.sum(axis=0)
    want[circ.nc] = full.any(axis=1).sum()
    want[circ.nc + 1] = total
    t0 = np.load(tmp_path / "tally0.npy")
    t1 = np.load(tmp_path / "tally1.npy")
    np.testing.assert_array_equal(t0, want)
    np.testing.assert_array_equal(t1, want)
    cat = np.concatenate([np.load(tmp_path / "out0.npy"), np.load(tmp_path / "out1.npy")])
    np.testing.assert_array_equal(cat, full)
