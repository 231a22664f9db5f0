#!/usr/bin/env python
"""Multi-GPU parity (run under torchrun, one rank per GPU): N GPUs == 1 GPU, bit for bit.

Every rank computes its ``shard_range`` of a sampler run and of an enumerator run on its own
GPU; rank 0 also computes the WHOLE range on its GPU.  Then
  * the all-reduced tallies (NCCL) must equal the whole-range tallies, entry by entry;
  * the ranks' packed outcome bits, gathered to rank 0 and concatenated in rank order, must
    equal the whole-range bits, word by word.
Prints one JSON line per case on rank 0 and exits non-zero on any mismatch.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port 29520 tools/verify_multi.py
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from qiskit_qec_b200 import _capi  # noqa: E402
from qiskit_qec_b200.circuits import (arc_circuit, heavy_hex_27_arc_layout,  # noqa: E402
                                      surface_code_circuit)
from qiskit_qec_b200.locations import tag_fault_locations  # noqa: E402
from qiskit_qec_b200.noise import standard_depolarizing_model  # noqa: E402
from qiskit_qec_b200.sharding import allreduce_tally, shard_range  # noqa: E402


def main():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    stream = torch.cuda.current_stream().cuda_stream
    ok_all = True

    def finish(name, total, nc, my_bits, my_lo, my_hi, tally, whole):
        """gather the shards' bits to rank 0, all-reduce the tallies, compare with `whole`"""
        nonlocal ok_all
        allreduce_tally(tally)
        nw_total = (total + 31) // 32
        ranges = [shard_range(total, r, world) for r in range(world)]
        if world > 1:
            # equal-size padded gather (shards differ by at most one lane word)
            maxw = max((hi - lo + 31) // 32 for lo, hi in ranges) or 1
            pad = torch.zeros((nc, maxw), dtype=torch.int32, device=dev)
            pad[:, :my_bits.shape[1]] = my_bits
            parts = [torch.zeros_like(pad) for _ in range(world)] if rank == 0 else None
            dist.gather(pad, parts, dst=0)
        else:
            parts = [my_bits]
        if rank == 0:
            w_bits, w_tally = whole
            cat = torch.cat([parts[r][:, :(hi - lo + 31) // 32] for r, (lo, hi) in enumerate(ranges)], dim=1)
            bits_ok = cat.shape[1] == nw_total and bool(torch.equal(cat, w_bits))
            tally_ok = bool(torch.equal(tally, w_tally))
            ok_all = ok_all and bits_ok and tally_ok
            print(json.dumps({"case": name, "n_gpus": world, "units": total, "bits_equal": bits_ok,
                              "tallies_equal": tally_ok, "words": nw_total,
                              "kernel": _capi.last_kernel()}), flush=True)

    def sampler_case(name, circ, n_shots, p, shot0):
        model = standard_depolarizing_model(p)
        idx, labels = tag_fault_locations(circ.labels, model)
        types = model.get_pauli_error_types()
        names = list(dict.fromkeys(labels))
        cls = {n: i for i, n in enumerate(names)}
        g = _capi.Circuit(circ.nq, circ.nc, circ.ops, device=local)
        sm = _capi.Sampler(g, idx, [cls[lab] for lab in labels],
                           [model.get_error_probability(n) for n in names],
                           [[_capi.pauli_code(pp) for pp in types[n]] for n in names],
                           [[model.get_pauli_weight(n, pp) for pp in types[n]] for n in names], 4242)
        lo, hi = shard_range(n_shots, rank, world)
        tl = _capi.tally_len(circ.nc)
        tally = torch.zeros(tl, dtype=torch.int64, device=dev)
        bits = torch.zeros((circ.nc, max((hi - lo + 31) // 32, 0)), dtype=torch.int32, device=dev)
        if hi > lo:
            sm.sample_dev(shot0 + lo, hi - lo, bits.data_ptr(), tally.data_ptr(), stream)
        whole = None
        if rank == 0:
            w_tally = torch.zeros(tl, dtype=torch.int64, device=dev)
            w_bits = torch.zeros((circ.nc, (n_shots + 31) // 32), dtype=torch.int32, device=dev)
            sm.sample_dev(shot0, n_shots, w_bits.data_ptr(), w_tally.data_ptr(), stream)
            whole = (w_bits, w_tally)
        torch.cuda.synchronize()
        sm.check()
        finish(name, n_shots, circ.nc, bits, lo, hi, tally, whole)
        sm.close(); g.close()

    def enum_case(name, circ, order, limit):
        model = standard_depolarizing_model()
        idx, labels = tag_fault_locations(circ.labels, model)
        types = model.get_pauli_error_types()
        g = _capi.Circuit(circ.nq, circ.nc, circ.ops, device=local)
        en = _capi.Enumerator(g, order, idx, [[_capi.pauli_code(pp) for pp in types[lab]] for lab in labels])
        total = min(en.num_paths(), limit)
        lo, hi = shard_range(total, rank, world)
        tl = _capi.tally_len(circ.nc)
        tally = torch.zeros(tl, dtype=torch.int64, device=dev)
        bits = torch.zeros((circ.nc, max((hi - lo + 31) // 32, 0)), dtype=torch.int32, device=dev)
        if hi > lo:
            en.range_dev(lo, hi, bits.data_ptr(), tally.data_ptr(), stream)
        whole = None
        if rank == 0:
            w_tally = torch.zeros(tl, dtype=torch.int64, device=dev)
            w_bits = torch.zeros((circ.nc, (total + 31) // 32), dtype=torch.int32, device=dev)
            en.range_dev(0, total, w_bits.data_ptr(), w_tally.data_ptr(), stream)
            whole = (w_bits, w_tally)
        torch.cuda.synchronize()
        finish(name, total, circ.nc, bits, lo, hi, tally, whole)
        en.close(); g.close()

    sampler_case("cfg2 sampler d=5 T=5 p=1e-3, 2e7 shots", surface_code_circuit(5, 5), 20_000_000, 1e-3, 64)
    sampler_case("cfg4 sampler d=11 T=11 p=1e-3, 4e5 shots (bits too)", surface_code_circuit(11, 11),
                 400_000 + 37, 1e-3, 0)
    enum_case("cfg3 enumerator d=5 T=5 order 2, all paths", surface_code_circuit(5, 5), 2, 1 << 62)
    links, sched, color = heavy_hex_27_arc_layout()
    enum_case("cfg5 enumerator ARC T=3 order 3, first 3e7 paths", arc_circuit(links, 3, "zx", sched, color),
              3, 30_000_000)
    if world > 1:
        flag = torch.tensor([1 if ok_all else 0], device=dev)
        dist.broadcast(flag, src=0)
        ok_all = bool(flag.item())
        dist.destroy_process_group()
    sys.exit(0 if ok_all else 1)


if __name__ == "__main__":
    main()
