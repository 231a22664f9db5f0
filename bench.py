#!/usr/bin/env python
"""bench.py -- BASELINE.json metric: FaultSampler shots/s (& FaultEnumerator faults/s) on the
d=5 surface-code memory circuit.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the hot path over one batch: FaultSampler on SurfaceCodeCircuit
d=5 T=5, p=1e-3 circuit-level noise, --shots shots (default 1e8 = BASELINE configs[1]) per
GPU, writing the packed outcome bits (1.81 GB, larger than L2) and the tallies.  N>1 is weak
scaling: every rank draws its own disjoint shot range and the tallies are all-reduced over
NCCL.  One JSON line on stdout (rank 0).
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "FaultSampler shots/s & FaultEnumerator faults/s, d=5 surface, 1-8 B200"
P_NOISE = 1e-3


def make_inputs(d=5, T=5):
    from qiskit_qec_b200.circuits import surface_code_circuit
    from qiskit_qec_b200.locations import tag_fault_locations
    from qiskit_qec_b200.noise import standard_depolarizing_model
    circ = surface_code_circuit(d, T)
    model = standard_depolarizing_model(P_NOISE)
    idx, labels = tag_fault_locations(circ.labels, model)
    types = model.get_pauli_error_types()
    return circ, model, idx, labels, types


# ---------------------------------------------------------------------------- reference arm
def _ref_worker(args):
    seed, n_shots, d, T = args
    from oracle import ref
    R = ref.module()
    circ, model, idx, labels, types = make_inputs(d, T)
    lw = {lab: [(p, model.get_pauli_weight(lab, p)) for p in types[lab]] for lab in set(labels)}
    lp = {lab: model.get_error_probability(lab) for lab in set(labels)}
    fs = R._CFaultSampler(circ.nq, circ.nc, circ.ops, idx, labels, lw, lp, seed)
    t0 = time.perf_counter()
    blk = fs.sample(n_shots)
    dt = time.perf_counter() - t0
    return len(blk), dt


def reference_rate(n_procs: int, shots_per_proc: int, d=5, T=5, seed0=0):
    """shots/s of the UNMODIFIED reference _CFaultSampler.sample (oracle/_ref) on n_procs host
    processes (the reference is single-threaded; this is its embarrassingly parallel bound)."""
    import multiprocessing as mp
    ctx = mp.get_context("fork")
    t0 = time.perf_counter()
    with ctx.Pool(n_procs) as pool:
        res = pool.map(_ref_worker, [(seed0 + i, shots_per_proc, d, T) for i in range(n_procs)])
    wall = time.perf_counter() - t0
    total = sum(r[0] for r in res)
    return total / max(r[1] for r in res), total, wall


def reference_enum_rate(n_paths: int = 100_000, d=5, T=5):
    """faults/s of the UNMODIFIED reference _CFaultEnumerator.enumerate (order 2) on ONE host core
    (the reference as shipped is single-threaded), over the first n_paths paths."""
    from oracle import ref
    R = ref.module()
    circ, model, idx, labels, types = make_inputs(d, T)
    fe = R._CFaultEnumerator(2, circ.nq, circ.nc, circ.ops, idx, labels, [types[lab] for lab in labels])
    t0 = time.perf_counter()
    n = 0
    while n < n_paths and not fe.done():
        n += len(fe.enumerate(10000))
    return n / (time.perf_counter() - t0), n


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import ref
    cores = host_cores()
    config = {"workload": "FaultSampler d=5 T=5 surface code, p=1e-3 circuit-level noise "
                          "(BASELINE configs[1]); bounded sample per step",
              "host_cores_used": cores}
    if not ref.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref not built"}))
        return
    per_proc = args.ref_shots
    for _ in range(args.warmup):
        reference_rate(cores, max(per_proc // 10, 100))
    rates, t0 = [], time.perf_counter()
    for s in range(args.steps):
        r, total, _ = reference_rate(cores, per_proc, seed0=1000 * s)
        rates.append(r)
    wall = time.perf_counter() - t0
    value = statistics.median(rates)
    # the reference as shipped: one process, one core; and its enumerator
    one_shots = max(per_proc // 2, 1000)
    one_core, _, _ = reference_rate(1, one_shots, seed0=77)
    enum_rate, enum_n = reference_enum_rate(100_000)
    line = {"metric": METRIC, "value": value, "unit": "shots/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * wall / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32",
            "data": "synthetic", "impl": "reference", "config": config,
            "cpu_baseline": {"value": value, "unit": "shots/s", "cores": cores, "kind": "reference",
                             "sample": f"{cores} processes x {per_proc} shots per step through "
                                       "reference _CFaultSampler.sample (oracle/_ref)",
                             "one_core": {"value": one_core, "unit": "shots/s", "cores": 1,
                                          "sample": f"1 process x {one_shots} shots: the reference as "
                                                    "shipped (single-threaded)"},
                             "enumerator_one_core": {
                                 "value": enum_rate, "unit": "faults/s", "cores": 1,
                                 "sample": f"first {enum_n} order-2 paths of d=5 T=5 through reference "
                                           "_CFaultEnumerator.enumerate"}},
            "e2e": {"value": value, "unit": "shots/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ---------------------------------------------------------------------------- clocks sampler
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.FIELDS}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ---------------------------------------------------------------------------- our arm
def run_ours(args):
    import torch
    import gc
    import torch.distributed as dist
    import numpy as np

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    cores = host_cores()

    from qiskit_qec_b200 import _capi
    from qiskit_qec_b200.sharding import allreduce_tally, shard_range
    if not torch.cuda.is_available() or _capi.device_count() < 1:
        raise RuntimeError("bench.py: no CUDA device; the product has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    circ, model, idx, labels, types = make_inputs()
    names = list(dict.fromkeys(labels))
    cls = {n: i for i, n in enumerate(names)}
    nc = circ.nc
    n_shots = args.shots
    n_words = (n_shots + 31) // 32

    # the host buffers the C ABI reads (qec_circuit_create: int32 row offsets + row data;
    # qec_sampler_create: int32 location tables + the class tables), laid out once: every
    # end-to-end step below hands THESE to the create calls and pays their host -> device copies
    csr_rows = _capi.rows_to_csr(circ.ops)
    loc_op32 = np.ascontiguousarray(idx, dtype=np.int32)
    loc_cls32 = np.asarray([cls[lab] for lab in labels], dtype=np.int32)
    class_p = [model.get_error_probability(n) for n in names]
    class_codes = [[_capi.pauli_code(p) for p in types[n]] for n in names]
    class_w = [[model.get_pauli_weight(n, p) for p in types[n]] for n in names]

    def new_sampler(seed=0):
        g = _capi.Circuit(circ.nq, nc, csr_rows, device=local)
        s = _capi.Sampler(g, loc_op32, loc_cls32, class_p, class_codes, class_w, seed)
        return g, s

    g, sm = new_sampler()
    w_bit = g.bit_ops
    bits = torch.empty((nc, n_words), dtype=torch.int32, device=dev)   # 1.81 GB at 1e8 shots
    # one tally buffer per step: the all-reduce of step i runs on a side stream while the
    # kernel of step i + 1 runs (the only collective of the path, 1.7 KB; the timed region ends
    # after the last all-reduce)
    tallies = [torch.zeros(_capi.tally_len(nc), dtype=torch.int64, device=dev)
               for _ in range(args.warmup + args.steps)]
    tally = tallies[-1]
    main = torch.cuda.current_stream()
    stream = main.cuda_stream
    side = torch.cuda.Stream(device=dev) if world > 1 else None

    def step(i):
        # a fresh, disjoint global shot range per (rank, step); tallies all-reduced over NCCL
        lo = ((rank * 1_000_003 + i) * n_words) * 32
        t_i = tallies[i]
        t_i.zero_()
        sm.sample_dev(lo, n_shots, bits.data_ptr(), t_i.data_ptr(), stream)
        if side is not None:
            side.wait_stream(main)
            with torch.cuda.stream(side):
                allreduce_tally(t_i)

    # clocks / throttle reasons are sampled (nvidia-smi, 100 ms period) from before the warm-up,
    # through the timed region and a second or so of the same load after it
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
        t_wait = time.perf_counter()
        while not clocks.rows and time.perf_counter() - t_wait < 3.0:
            time.sleep(0.05)
    for i in range(args.warmup):
        step(i)
    if side is not None:
        main.wait_stream(side)
    torch.cuda.synchronize()
    sm.check()
    if world > 1:
        dist.barrier()  # every rank enters the timed region together
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    torch.cuda.synchronize()
    ev[0].record()
    for i in range(args.steps):
        step(args.warmup + i)
        if side is not None and i == args.steps - 1:
            main.wait_stream(side)  # the timed region covers every all-reduce
        ev[i + 1].record()
    torch.cuda.synchronize()
    sm.check()
    sampler_kernel = _capi.last_kernel()
    if world > 1:
        dist.barrier()
    total_ms = ev[0].elapsed_time(ev[-1])
    step_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    value = world * n_shots * args.steps / (total_ms * 1e-3)
    assert int(tally[nc + 1].item()) == n_shots * world

    # kernel-only launch time of the dominant kernel (k_sample), CUDA events inside the library
    kern_ms = statistics.mean(step_ms)

    # The timed region lasts a few ms, less than one period of the clocks sampler: keep the same
    # load running (untimed steps of the same kernel) until a dozen samples are in, then stop the
    # sampler -- an nvidia-smi query every 100 ms takes the driver's lock and showed up as +1.5 ms on
    # single steps of the API-heavy end-to-end legs below, which are PCIe-bound and need no clocks.
    if rank == 0:
        t_load = time.perf_counter()
        n0 = len(clocks.rows)
        while len(clocks.rows) < n0 + 12 and time.perf_counter() - t_load < 2.5:
            for i in range(32):  # (no collective here: only rank 0 runs this loop)
                sm.sample_dev(((rank * 1_000_003 + i) * n_words) * 32, n_shots, bits.data_ptr(),
                              tallies[0].data_ptr(), stream)
            torch.cuda.synchronize()
    clk = clocks.stop() if rank == 0 else None
    if side is not None:
        main.wait_stream(side)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()

    # ---- end to end: the public calls with HOST buffers, copies inside the timed region
    host_tally = torch.empty(_capi.tally_len(nc), dtype=torch.int64, pin_memory=True)
    tab_bytes = int(sm.tables().nbytes) + 4 * len(circ.ops)
    del bits
    torch.cuda.empty_cache()

    def settle(fn, max_iter=12):
        """untimed passes until three in a row are within 5 % of each other AND below 1.5x the
        best seen: on a fresh box the first passes over a newly pinned buffer run 3-6x slower,
        and two equally slow passes must not end the warm-up"""
        seen = []
        for _ in range(max_iter):
            if world > 1:
                dist.barrier()  # (all ranks pass together, as in the timed steps: they share the host's memory system)
            ts = time.perf_counter()
            fn(0)
            el = torch.tensor([time.perf_counter() - ts], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(el, op=dist.ReduceOp.MAX)  # one decision for all ranks
            seen.append(float(el.item()))
            last = seen[-3:]
            if len(seen) >= 3 and max(last) <= 1.05 * min(last) and max(last) <= 1.5 * min(seen):
                break
        return seen

    def timed_e2e(fn, steps):
        warm = settle(fn)
        # host-timed steps: keep the cyclic collector out of them, as timeit does (a generation-2
        # pass of a process with torch loaded takes tens of ms -- several 4 ms steps)
        gc.collect()
        gc_was = gc.isenabled()
        gc.disable()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        times = []
        for i in range(steps):
            ts = time.perf_counter()
            fn(1 + i)
            times.append(time.perf_counter() - ts)
        elapsed = time.perf_counter() - t0
        if gc_was:
            gc.enable()
        t = torch.tensor([elapsed], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return world * n_shots * steps / float(t.item()), times, warm

    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    # (1) compact call -- the headline: the sampled faults of every shot + tallies.  Lossless:
    # labels and errors ARE the list, the outcome is the XOR of the faults' signatures; about
    # 1.6 B/shot cross PCIe instead of 18.1
    # (two bytes per fault on the wire where the lane-stream kernel serves the sampler: delta-coded
    # entries in slot order, qec_sampler_sample_events16; else three: qec_sampler_sample_events24)
    ev16 = "lane-streams" in sampler_kernel
    ev_cap = int(n_shots * max(1.2 * len(idx) * P_NOISE, 0.05)) * (5 if ev16 else 4) // 4 + 65536
    ev_pin = torch.empty(ev_cap * (2 if ev16 else 3), dtype=torch.uint8, pin_memory=True)
    wc_pin = torch.empty(n_words, dtype=torch.int16, pin_memory=True)
    pinned = (wc_pin.numpy().view(np.uint16), ev_pin.numpy().view(np.uint16) if ev16 else ev_pin.numpy())
    ev_bytes = [0]

    def e2e_compact(i):
        g2, s2 = new_sampler()
        wc, ev, tl = s2.sample_events(((rank * 1_000_003 + 500_000 + i) * n_words) * 32, n_shots,
                                      want_tally=True, pinned=pinned, packed24=not ev16, packed16=ev16)
        ev_bytes[0] = int(ev.nbytes + wc.nbytes + tl.nbytes)
        assert int(tl[nc + 1]) == n_shots
        s2.close()
        g2.close()

    e2e_value, e2e_times, e2e_warm = timed_e2e(e2e_compact, e2e_steps)
    compact_kernel = _capi.last_kernel()
    del ev_pin, wc_pin, pinned

    # (2) dense call: packed outcome bits of every shot (18.1 B/shot) + tallies, pinned memory
    host_bits = torch.empty((nc, n_words), dtype=torch.int32, pin_memory=True)

    def e2e_dense(i):
        g2, s2 = new_sampler()
        s2.sample_host_ptr(((rank * 1_000_003 + 700_000 + i) * n_words) * 32, n_shots,
                           host_bits.data_ptr(), host_tally.data_ptr())
        s2.close()
        g2.close()

    dense_value, dense_times, dense_warm = timed_e2e(e2e_dense, e2e_steps)
    assert int(host_tally[nc + 1].item()) == n_shots
    dense_bytes = int(host_bits.numel() * 4 + host_tally.numel() * 8)
    del host_bits

    # (3) the literal plug-in call of the reference's API: _CFaultSampler(...).sample(N) -> N
    # Python tuples (index, labels, errors, outcome list); building the objects dominates
    plugin = None
    if rank == 0 and args.plugin_shots > 0:
        from qiskit_qec_b200.analysis.extensions import _CFaultSampler
        lw = {lab: [(p, model.get_pauli_weight(lab, p)) for p in types[lab]] for lab in set(labels)}
        lp = {lab: model.get_error_probability(lab) for lab in set(labels)}
        fs = _CFaultSampler(circ.nq, nc, circ.ops, idx, labels, lw, lp, 5)
        fs.sample(10_000)
        ts = time.perf_counter()
        blk = fs.sample(args.plugin_shots)
        dt = time.perf_counter() - ts
        assert len(blk) == args.plugin_shots and blk[0][0] == -1 and len(blk[0][3]) == nc
        plugin = {"value": args.plugin_shots / dt, "unit": "shots/s", "shots": args.plugin_shots,
                  "ms": 1e3 * dt,
                  "call": "_c_analysis._CFaultSampler(...).sample(N): the reference's own signature "
                          "and return type (list of N tuples); GPU sampling + host tuple building"}
        del blk, fs

    # secondary metric: FaultEnumerator order 2 on the same circuit (BASELINE configs[2]), index
    # range sharded evenly across ranks, outcomes + tallies on device.  The tally all-reduce of a
    # step runs on a side stream under the next step's kernel, as the sampler's does (round 1 ran
    # it on the main stream: 60 us of launch + NCCL latency against 36 us of kernel at 8 GPUs)
    paulis = [types[lab] for lab in labels]
    enum = None
    if not args.no_enum:
        reps = 8
        en = _capi.Enumerator(g, 2, idx, [[_capi.pauli_code(p) for p in pl] for pl in paulis])
        total = en.num_paths()
        lo, hi = shard_range(total, rank, world)
        ebits = torch.empty((nc, (hi - lo + 31) // 32), dtype=torch.int32, device=dev)
        etal = [torch.zeros(_capi.tally_len(nc), dtype=torch.int64, device=dev) for _ in range(reps + 3)]

        def estep(i):
            en.range_dev(lo, hi, ebits.data_ptr(), etal[i].data_ptr(), stream)
            if side is not None:
                side.wait_stream(main)
                with torch.cuda.stream(side):
                    allreduce_tally(etal[i])

        for i in range(3):
            estep(i)
        if side is not None:
            main.wait_stream(side)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(reps):
            estep(3 + i)
        if side is not None:
            main.wait_stream(side)  # the timed region covers every all-reduce
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ems = float(t.item()) / reps
        assert int(etal[-1][nc + 1].item()) == total
        enum = {"metric": "FaultEnumerator order-2 faults/s", "value": total / (ems * 1e-3),
                "unit": "faults/s", "paths": total, "ms": ems, "scaling": "strong",
                "workload": "order=2 exhaustive, SurfaceCodeCircuit d=5 T=5 (BASELINE configs[2])",
                "kernel": _capi.last_kernel(),
                "hbm_frac": (total * nc / 8.0 / (ems * 1e-3) / 1e9 / peaks_hbm()) if world == 1 else None}
        en.close()
        del ebits
        # the same circuit at order 3 (4.8e10 paths; the first 4e9, tallies only): the strong-scaling
        # point that is not launch-bound (order 2 is 0.03 ms of kernel per rank at 8 GPUs)
        en3 = _capi.Enumerator(g, 3, idx, [[_capi.pauli_code(p) for p in pl] for pl in paulis])
        total3 = min(en3.num_paths(), 4_000_000_000)
        lo3, hi3 = shard_range(total3, rank, world)
        t3 = [torch.zeros(_capi.tally_len(nc), dtype=torch.int64, device=dev) for _ in range(4)]

        def estep3(i):
            en3.range_dev(lo3, hi3, 0, t3[i].data_ptr(), stream)
            if side is not None:
                side.wait_stream(main)
                with torch.cuda.stream(side):
                    allreduce_tally(t3[i])

        estep3(0)
        if side is not None:
            main.wait_stream(side)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(3):
            estep3(1 + i)
        if side is not None:
            main.wait_stream(side)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ems3 = float(t.item()) / 3
        assert int(t3[-1][nc + 1].item()) == total3
        enum["order3"] = {"metric": "FaultEnumerator order-3 faults/s (tallies only)", "value": total3 / (ems3 * 1e-3),
                          "unit": "faults/s", "paths": total3, "ms": ems3, "scaling": "strong",
                          "workload": "order=3, first 4e9 of 4.76e10 paths, SurfaceCodeCircuit d=5 T=5",
                          "kernel": _capi.last_kernel()}
        en3.close()

    # the gate-stream (frame-propagation) kernels the north star names: the same workloads under
    # QEC_B200_SAMPLER=frame / QEC_B200_ENUM=frame.  These EXECUTE the W_bit/32 lane-ops per shot,
    # so their integer-pipe roofline fraction is a real one (filled in below, against k_int_peak)
    frame = None
    if rank == 0 and world == 1 and not args.no_frame:
        os.environ["QEC_B200_SAMPLER"] = "frame"
        os.environ["QEC_B200_ENUM"] = "frame"
        try:
            fshots = min(n_shots, args.frame_shots)
            fwords = (fshots + 31) // 32
            gf, sf = new_sampler(seed=1)
            fbits = torch.empty((nc, fwords), dtype=torch.int32, device=dev)
            ftal = torch.zeros(_capi.tally_len(nc), dtype=torch.int64, device=dev)
            f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            for i in range(3):
                sf.sample_dev(i * fwords * 32, fshots, fbits.data_ptr(), ftal.data_ptr(), stream)
            torch.cuda.synchronize()
            sf.check()
            f0.record()
            for i in range(5):
                sf.sample_dev((3 + i) * fwords * 32, fshots, fbits.data_ptr(), ftal.data_ptr(), stream)
            f1.record()
            torch.cuda.synchronize()
            sf.check()
            fms, fkern = f0.elapsed_time(f1) / 5, _capi.last_kernel()
            sf.close()
            del fbits
            en_f = _capi.Enumerator(gf, 2, idx, [[_capi.pauli_code(p) for p in pl] for pl in paulis])
            tot_f = en_f.num_paths()
            ebf = torch.empty((nc, (tot_f + 31) // 32), dtype=torch.int32, device=dev)
            for _ in range(2):
                en_f.range_dev(0, tot_f, ebf.data_ptr(), ftal.data_ptr(), stream)
            torch.cuda.synchronize()
            f0.record()
            for _ in range(5):
                en_f.range_dev(0, tot_f, ebf.data_ptr(), ftal.data_ptr(), stream)
            f1.record()
            torch.cuda.synchronize()
            frame = {"sampler": {"value": fshots / (fms * 1e-3), "unit": "shots/s", "units": fshots,
                                 "ms": fms, "kernel": fkern},
                     "enumerator": {"value": tot_f / (f0.elapsed_time(f1) / 5 * 1e-3), "unit": "faults/s",
                                    "units": tot_f, "ms": f0.elapsed_time(f1) / 5,
                                    "kernel": _capi.last_kernel()}}
            en_f.close()
            gf.close()
            del ebf
        finally:
            os.environ.pop("QEC_B200_SAMPLER", None)
            os.environ.pop("QEC_B200_ENUM", None)

    # CPU baseline (rank 0, N=1 only): the unmodified reference in a child process, after
    # the GPU measurements so that its worker processes cannot disturb them
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            out = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference",
                                  "--steps", "1", "--warmup", "0", "--ref-shots",
                                  str(args.ref_shots)], capture_output=True, text=True,
                                 timeout=900).stdout.strip().splitlines()
            ref = json.loads(out[-1]) if out else {}
            cpu_baseline = ref.get("cpu_baseline")
        except (subprocess.SubprocessError, ValueError, OSError):
            cpu_baseline = None

    if rank == 0:
        hbm_peak, peak_src = peaks_hbm(), peaks_src()
        int_peak, clk_mhz = _capi.measure_int_peak(local, 40.0)
        # dram bytes per shot of the dominant kernel, from this round's committed ncu --set full
        # capture -- used only when that capture is of the kernel variant that just ran, so that a
        # changed kernel cannot silently inherit a stale figure (smaller 12)
        traffic, traffic_src = None, None
        try:
            cap = json.load(open(os.path.join(ROOT, "profiles", "r02", "ncu_dominant_kernel.json")))
            if cap.get("variant_key") and cap["variant_key"] in sampler_kernel:
                traffic = float(cap["dram_bytes_per_shot"]) * n_shots
                traffic_src = ("profiles/r02/ncu_dominant_kernel.json: dram read+write bytes per shot x "
                               "shots per launch, capture of the same kernel variant")
        except (OSError, ValueError, KeyError):
            pass
        achieved = n_shots * nc / 8.0 / (kern_ms * 1e-3) / 1e9
        lane_ops = w_bit / 32.0
        if frame:
            for k in ("sampler", "enumerator"):
                ach = frame[k]["units"] * lane_ops / (frame[k]["ms"] * 1e-3)
                frame[k]["int_roofline"] = {"bound": "int32-alu", "achieved": ach, "peak": int_peak,
                                            "unit": "lane-ops/s", "frac": ach / int_peak,
                                            "lane_ops_per_unit": lane_ops}
        line = {
            "metric": METRIC, "value": value, "unit": "shots/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32",
            "data": "synthetic",
            "config": {"workload": f"FaultSampler {n_shots} shots/GPU/step, SurfaceCodeCircuit d=5 "
                                   "T=5, p=1e-3 circuit-level noise (BASELINE configs[1]); packed "
                                   "outcomes + tallies written per step",
                       "l2": "output per step (1.81 GB at 1e8 shots) is larger than L2; no reuse",
                       "timing": "CUDA events on the launching stream, max over ranks",
                       "host_cores": cores},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                         "frac": achieved / hbm_peak, "traffic": traffic,
                         "traffic_source": traffic_src, "kernel": sampler_kernel,
                         "peak_source": peak_src, "algorithmic_bytes_per_shot": nc / 8.0},
            "int_peak": {"value": int_peak, "unit": "lane-ops/s", "sm_clock_mhz": clk_mhz,
                         "source": "k_int_peak: 8 independent chains of inline-PTX lop3.b32 per thread, "
                                   "counted ops == executed LOP3 (profiles/r02/sass_k_int_peak.txt), "
                                   "measured in this run",
                         "note": "the integer-pipe roofline applies to the gate-stream kernels "
                                 "(frame.*.int_roofline); the default linear-response kernels do not "
                                 "execute the gate stream and are judged on the HBM roofline only"},
            "e2e": {"value": e2e_value, "unit": "shots/s", "h2d_bytes_per_step": tab_bytes,
                    "d2h_bytes_per_step": ev_bytes[0], "steps": e2e_steps,
                    "ms_per_step": [round(x * 1e3, 2) for x in e2e_times],
                    "warmup_ms": [round(x * 1e3, 1) for x in e2e_warm],
                    "call": "qec_circuit_create + qec_sampler_create (host int32 / double arrays as the C ABI "
                            "takes them, laid out before the timed region) + " +
                            ("qec_sampler_sample_events16 (pinned host buffers: per-word counts + the sampled "
                             "faults of every shot, delta-coded in two bytes each" if ev16 else
                             "qec_sampler_sample_events24 (pinned host buffers: per-word counts + the sampled "
                             "faults of every shot, three bytes each") +
                            ", + tallies) + destroy; lossless compact form of the shots",
                    "kernel": compact_kernel},
            "e2e_dense": {"value": dense_value, "unit": "shots/s", "h2d_bytes_per_step": tab_bytes,
                          "d2h_bytes_per_step": dense_bytes,
                          "ms_per_step": [round(x * 1e3, 2) for x in dense_times],
                          "warmup_ms": [round(x * 1e3, 1) for x in dense_warm],
                          "call": "qec_circuit_create + qec_sampler_create + qec_sampler_sample (host "
                                  "pointers, pinned: packed outcome bits of every shot + tallies) + destroy"},
            "e2e_plugin": plugin,
            "gpu_launches": args.steps, "clocks": clk, "cpu_baseline": cpu_baseline,
            "enumerator": enum, "frame": frame, "kernel_ms_per_launch": kern_ms,
            "step_ms": [round(x, 4) for x in step_ms],
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def _peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        return {}


def peaks_hbm() -> float:
    # fallback: the B200 HBM figure of /opt/skills/guides/B200_PROFILING.md
    return float(_peaks().get("hbm_gbs", 6650.0))


def peaks_src() -> str:
    return "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in _peaks() else "fallback (B200_PROFILING.md)"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--shots", type=int, default=100_000_000, help="shots per GPU per step")
    ap.add_argument("--ref-shots", type=int, default=40_000,
                    help="reference shots per host process per step (bounded CPU sample)")
    ap.add_argument("--e2e-steps", type=int, default=8,
                    help="timed steps of each end-to-end leg (at most --steps): a single slow step -- on an "
                         "8-GPU box the ranks share one host memory system -- weighs 1/8, not 1/3")
    ap.add_argument("--plugin-shots", type=int, default=1_000_000,
                    help="shots of the literal _CFaultSampler.sample(N) call (0: skip)")
    ap.add_argument("--frame-shots", type=int, default=50_000_000,
                    help="shots per launch of the gate-stream (frame) sampler sub-record")
    ap.add_argument("--no-frame", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-enum", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
