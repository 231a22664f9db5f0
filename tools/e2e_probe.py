"""Where the end-to-end leg of bench.py spends its time: sampler creation, the host-pointer
sample call (chunked device->host streaming) and destruction, per step.

  [QEC_B200_PIPE_MB=N] python tools/e2e_probe.py
"""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from qiskit_qec_b200 import _capi
circ, model, idx, labels, types = bench.make_inputs()
names = list(dict.fromkeys(labels)); cls = {n: i for i, n in enumerate(names)}
nc = circ.nc; n_shots = 100_000_000; n_words = n_shots // 32
host_bits = torch.empty((nc, n_words), dtype=torch.int32, pin_memory=True)
host_tally = torch.empty(_capi.tally_len(nc), dtype=torch.int64, pin_memory=True)
def new():
    g = _capi.Circuit(circ.nq, nc, circ.ops, device=0)
    s = _capi.Sampler(g, idx, [cls[l] for l in labels], [model.get_error_probability(n) for n in names],
                      [[_capi.pauli_code(p) for p in types[n]] for n in names],
                      [[model.get_pauli_weight(n, p) for p in types[n]] for n in names], 0)
    return g, s
for mb in (os.environ.get("QEC_B200_PIPE_MB", "default"),):
    for i in range(8):
        t0 = time.perf_counter(); g, s = new(); t1 = time.perf_counter()
        s.sample_host_ptr(i * n_shots, n_shots, host_bits.data_ptr(), host_tally.data_ptr()); t2 = time.perf_counter()
        s.close(); g.close(); t3 = time.perf_counter()
        print(f"pipe={mb} create {1e3*(t1-t0):.2f} ms  sample {1e3*(t2-t1):.2f} ms  close {1e3*(t3-t2):.2f} ms", flush=True)

# compact forms: fault events (pinned host buffers) and the K4 counts table
import numpy as np
ev_pin = torch.empty(int(n_shots * 0.9), dtype=torch.int32, pin_memory=True)
wc_pin = torch.empty(n_words, dtype=torch.int16, pin_memory=True)
pin = (wc_pin.numpy().view(np.uint16), ev_pin.numpy().view(np.uint32))
g, s = new()
for i in range(6):
    t0 = time.perf_counter()
    wc, ev, tl = s.sample_events(i * n_shots, n_shots, want_tally=True, pinned=pin)
    t1 = time.perf_counter()
    print(f"events: {1e3*(t1-t0):.2f} ms  n_events={len(ev)}  bytes={len(ev)*4 + len(wc)*2}  {_capi.last_kernel()[:40]}", flush=True)
for n in (10_000_000, 100_000_000):
    for i in range(3):
        t0 = time.perf_counter()
        rows, counts, nz, _ = s.counts(i * n, n)
        t1 = time.perf_counter()
        print(f"counts n={n}: {1e3*(t1-t0):.2f} ms  distinct={len(rows)} zero={nz}", flush=True)
s.close(); g.close()
