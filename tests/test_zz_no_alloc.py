"""qec_outcome_counts_noalloc_dev: K4 with its hash table in a fixed device array of the library
(no cudaMalloc / cudaFree inside the call), against the existing entry point
qec_outcome_histogram_rows_dev with a caller-provided table, on the same packed outcomes.

Integer work: the two must agree exactly -- the same set of distinct outcomes, the same
multiplicities, the same all-zero count.  Two calls in a row on different inputs check that the
shared array is cleared between calls.  Everything goes through the C ABI (device memory from
qec_dev_alloc, no torch), so the file passes alone in a fresh process and after the other tests.
This file sorts last on purpose.
"""

import ctypes as C

import numpy as np
import pytest

from helpers import make_sampler, sampler_inputs
from qiskit_qec_b200.circuits import surface_code_circuit
from qiskit_qec_b200.noise import standard_depolarizing_model

pytestmark = pytest.mark.gpu


class _Dev:
    """device buffers through the C ABI's own helpers"""

    def __init__(self, capi):
        self.capi, self.lib, self.ptrs = capi, capi.lib(), []

    def alloc(self, nbytes):
        p = C.c_void_p()
        self.capi._check(self.lib.qec_dev_alloc(C.c_int(0), C.c_size_t(max(nbytes, 8)), C.byref(p)))
        self.ptrs.append(p)
        return p

    def zero(self, p, nbytes):
        self.capi._check(self.lib.qec_dev_memset(C.c_int(0), p, C.c_int(0), C.c_size_t(nbytes), None))

    def fetch(self, p, dtype, count):
        out = np.zeros(count, dtype=dtype)
        self.capi._check(self.lib.qec_dev_to_host(C.c_int(0), out.ctypes.data_as(C.c_void_p), p,
                                                  C.c_size_t(out.nbytes), None))
        self.capi._check(self.lib.qec_stream_sync(C.c_int(0), None))
        return out

    def close(self):
        self.capi._check(self.lib.qec_stream_sync(C.c_int(0), None))
        for p in self.ptrs:
            self.lib.qec_dev_free(C.c_int(0), p)
        self.ptrs = []


def _sorted(rows, counts):
    order = np.lexsort(rows.T[::-1]) if len(rows) else np.zeros(0, dtype=np.int64)
    return rows[order], counts[order]


def test_noalloc_counts_equal_the_table_entry_point_twice_in_a_row(capi):
    circ = surface_code_circuit(3, 2)
    nc, rw = circ.nc, (circ.nc + 31) // 32
    idx, labels, lw, lp = sampler_inputs(circ, standard_depolarizing_model(0.01))
    g = capi.Circuit(circ.nq, nc, circ.ops)
    sm, _ = make_sampler(capi, g, idx, labels, lw, lp, seed=123)
    dev = _Dev(capi)
    try:
        cap, max_out = 1 << 16, 1 << 16
        keys, counts = dev.alloc(cap * 8), dev.alloc(cap * 8)
        rows, status = dev.alloc(cap * rw * 4), dev.alloc(3 * 8)
        o_counts, o_rows, o_status = dev.alloc(max_out * 8), dev.alloc(max_out * rw * 4), dev.alloc(4 * 8)
        seen = []
        # two calls in a row: different shots, different sizes, the second one ragged
        for lo, n in ((0, 60_000), (4096 * 40, 35_003)):
            nw = (n + 31) // 32
            bits = dev.alloc(nc * nw * 4)
            sm.sample_dev(lo, n, bits.value, 0)
            sm.check()
            # existing entry point, caller-provided table
            for p, nb in ((keys, cap * 8), (counts, cap * 8), (status, 24)):
                dev.zero(p, nb)
            capi.outcome_histogram_dev(g, bits.value, n, keys.value, counts.value, cap,
                                       d_rows=rows.value, d_status=status.value, n_words=nw)
            k = dev.fetch(keys, np.uint64, cap)
            occ = k != 0
            want_rows = dev.fetch(rows, np.uint32, cap * rw).reshape(cap, rw)[occ]
            want_cnt = dev.fetch(counts, np.uint64, cap)[occ]
            want_st = dev.fetch(status, np.uint64, 3)
            # the new entry point: no table passed, nothing allocated inside
            capi.outcome_counts_noalloc_dev(g, bits.value, n, nw, o_counts.value, o_rows.value, max_out,
                                            o_status.value)
            st = dev.fetch(o_status, np.uint64, 4)
            m = int(st[3])
            got_cnt = dev.fetch(o_counts, np.uint64, max_out)[:m]
            got_rows = dev.fetch(o_rows, np.uint32, max_out * rw).reshape(max_out, rw)[:m]
            assert [int(x) for x in st[:3]] == [int(x) for x in want_st], (st, want_st)
            assert m == int(want_st[0]) == len(want_rows) and int(st[1]) == 0
            assert int(got_cnt.sum()) + int(st[2]) == n
            a_rows, a_cnt = _sorted(got_rows, got_cnt)
            b_rows, b_cnt = _sorted(want_rows, want_cnt)
            assert np.array_equal(a_rows, b_rows) and np.array_equal(a_cnt, b_cnt)
            seen.append(m)
        assert seen[0] != seen[1] and min(seen) > 100   # two genuinely different tables
        # max_out smaller than the number of outcomes: bounded writes, the total still reported
        small = 50
        capi.outcome_counts_noalloc_dev(g, bits.value, n, nw, o_counts.value, o_rows.value, small,
                                        o_status.value)
        st = dev.fetch(o_status, np.uint64, 4)
        assert int(st[0]) == seen[1] and int(st[3]) == small
    finally:
        dev.close()
        sm.close()
        g.close()


def test_noalloc_rejects_sizes_beyond_its_fixed_table(capi):
    big = surface_code_circuit(11, 3)      # 481 clbits > 256
    g = capi.Circuit(big.nq, big.nc, big.ops)
    dev = _Dev(capi)
    try:
        buf = dev.alloc(1024)
        with pytest.raises(RuntimeError, match="at most 256 clbits"):
            capi.outcome_counts_noalloc_dev(g, buf.value, 32, 1, buf.value, buf.value, 1, buf.value)
    finally:
        dev.close()
        g.close()
