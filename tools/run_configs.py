#!/usr/bin/env python
"""Runs the five BASELINE.json configs through the GPU path (kernel-scope timing, outcomes and
tallies left on the device) and prints one JSON line each.  Sizes that do not fit a short
run are scaled down and say so.  Under torchrun every rank takes its shard
(qiskit_qec_b200.sharding) and the tallies are all-reduced.

  python tools/run_configs.py [--shots4 N] [--paths5 N]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from qiskit_qec_b200 import _capi  # noqa: E402
from qiskit_qec_b200.circuits import (arc_circuit, heavy_hex_27_arc_layout,  # noqa: E402
                                      repetition_code_circuit, surface_code_circuit)
from qiskit_qec_b200.locations import tag_fault_locations  # noqa: E402
from qiskit_qec_b200.noise import standard_depolarizing_model  # noqa: E402
from qiskit_qec_b200.sharding import allreduce_tally, shard_range  # noqa: E402


def setup(circ, p=1e-3):
    model = standard_depolarizing_model(p)
    idx, labels = tag_fault_locations(circ.labels, model)
    types = model.get_pauli_error_types()
    return model, idx, labels, types


def timed(fn, reps, dev, world):
    fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shots2", type=int, default=100_000_000)
    ap.add_argument("--shots4", type=int, default=1_000_000_000, help="global shots of config 4 (BASELINE: 1e9)")
    ap.add_argument("--paths5", type=int, default=400_000_000, help="order-3 paths of config 5 to run (all 3.4e8 by default)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    stream = torch.cuda.current_stream().cuda_stream

    def enum_cfg(name, circ, order, limit=None, bits=True):
        model, idx, labels, types = setup(circ)
        g = _capi.Circuit(circ.nq, circ.nc, circ.ops, device=local)
        paulis = [types[lab] for lab in labels]
        en = _capi.Enumerator(g, order, idx, [[_capi.pauli_code(p) for p in pl] for pl in paulis])
        total = en.num_paths()
        n = min(total, limit) if limit else total
        lo, hi = shard_range(n, rank, world)
        tally = torch.zeros(_capi.tally_len(circ.nc), dtype=torch.int64, device=dev)
        ebits = torch.empty((circ.nc, max((hi - lo + 31) // 32, 1)), dtype=torch.int32, device=dev) if bits else None

        def run():
            tally.zero_()
            if hi > lo:
                en.range_dev(lo, hi, ebits.data_ptr() if bits else 0, tally.data_ptr(), stream)
            allreduce_tally(tally)

        ms = timed(run, 3, dev, world)
        assert int(tally[circ.nc + 1].item()) == n
        if rank == 0:
            print(json.dumps({"config": name, "kind": "enumerator", "order": order, "paths_total": total,
                              "paths_run": n, "n_gpus": world, "ms": ms, "faults_per_s": n / ms * 1e3,
                              "nq": circ.nq, "nc": circ.nc, "n_ops": len(circ.ops), "locations": len(idx),
                              "packed_outcomes": bits, "kernel": _capi.last_kernel()}), flush=True)
        en.close(); g.close()

    def samp_cfg(name, circ, n_shots, bits=True):
        model, idx, labels, types = setup(circ)
        names = list(dict.fromkeys(labels))
        cls = {n: i for i, n in enumerate(names)}
        g = _capi.Circuit(circ.nq, circ.nc, circ.ops, device=local)
        sm = _capi.Sampler(g, idx, [cls[lab] for lab in labels],
                           [model.get_error_probability(n) for n in names],
                           [[_capi.pauli_code(p) for p in types[n]] for n in names],
                           [[model.get_pauli_weight(n, p) for p in types[n]] for n in names], 0)
        lo, hi = shard_range(n_shots, rank, world)
        tally = torch.zeros(_capi.tally_len(circ.nc), dtype=torch.int64, device=dev)
        sbits = torch.empty((circ.nc, max((hi - lo + 31) // 32, 1)), dtype=torch.int32, device=dev) if bits else None

        def run():
            tally.zero_()
            if hi > lo:
                sm.sample_dev(lo, hi - lo, sbits.data_ptr() if bits else 0, tally.data_ptr(), stream)
            allreduce_tally(tally)

        ms = timed(run, 3, dev, world)
        sm.check()
        assert int(tally[circ.nc + 1].item()) == n_shots
        if rank == 0:
            print(json.dumps({"config": name, "kind": "sampler", "shots": n_shots, "n_gpus": world, "ms": ms,
                              "shots_per_s": n_shots / ms * 1e3, "nq": circ.nq, "nc": circ.nc,
                              "n_ops": len(circ.ops), "locations": len(idx), "packed_outcomes": bits,
                              "mean_faults_per_shot": float((tally[circ.nc + 2:] *
                                                             torch.arange(64, device=dev)).sum() / n_shots),
                              "kernel": _capi.last_kernel()}), flush=True)
        sm.close(); g.close()

    enum_cfg("cfg1 rep d=3 T=2 order 1", repetition_code_circuit(3, 2), 1)
    samp_cfg("cfg2 surface d=5 T=5 sampler", surface_code_circuit(5, 5), args.shots2)
    enum_cfg("cfg3 surface d=5 T=5 order 2", surface_code_circuit(5, 5), 2)
    samp_cfg(f"cfg4 surface d=11 T=11 sampler, tallies only ({args.shots4} shots; BASELINE 1e9)",
             surface_code_circuit(11, 11), args.shots4, bits=False)
    links, schedule, color = heavy_hex_27_arc_layout()
    enum_cfg(f"cfg5 heavy-hex ARC T=3 order 3, tallies only (first {args.paths5} paths)",
             arc_circuit(links, 3, "zx", schedule, color), 3, limit=args.paths5, bits=False)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
