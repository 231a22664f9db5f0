"""ctypes binding of ``libqecb200.so`` -- the C ABI declared in ``include/qec_b200.h``.

Thin by design: numpy arrays in, numpy arrays out, one Python method per C entry point.
There is no fallback: if the shared library is missing or no CUDA device is usable the
calls raise.  Status codes map to the exception types pybind11 gives the reference's C++
exceptions (``std::logic_error`` -> RuntimeError, ``std::invalid_argument`` -> ValueError;
SURVEY.md 8(b) "Error conventions").
"""

from __future__ import annotations

import ctypes as C
import itertools
import os
from typing import Optional, Sequence, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libqecb200.so")

QEC_ERR_LOGIC, QEC_ERR_INVALID, QEC_ERR_CUDA, QEC_ERR_LIMIT, QEC_ERR_OVERFLOW = -1, -2, -3, -4, -5
TALLY_HIST_BINS = 64

# every symbol include/qec_b200.h declares (checked by tests/test_capi_symbols.py)
SYMBOLS = [
    "qec_last_error", "qec_version", "qec_device_count", "qec_measure_int_peak",
    "qec_circuit_create", "qec_circuit_destroy", "qec_circuit_sizes", "qec_circuit_bit_ops",
    "qec_propagate_batch",
    "qec_enumerator_create", "qec_enumerator_destroy", "qec_enumerator_num_paths",
    "qec_enumerator_unrank", "qec_enumerator_subset_rank", "qec_enumerator_range",
    "qec_enumerator_range_dev",
    "qec_sampler_create", "qec_sampler_destroy", "qec_sampler_sample", "qec_sampler_sample_dev",
    "qec_sampler_check", "qec_sampler_tables", "qec_sampler_variant", "qec_sampler_counts",
    "qec_sampler_sample_events",
    "qec_multi_sampler_create", "qec_multi_sampler_destroy", "qec_multi_sampler_sample",
    "qec_multi_sampler_tally_dev", "qec_multi_enumerator_create", "qec_multi_enumerator_destroy",
    "qec_multi_enumerator_num_paths", "qec_multi_enumerator_range", "qec_multi_enumerator_tally_dev",
    "qec_outcome_histogram_dev", "qec_outcome_histogram_rows_dev", "qec_outcome_counts_noalloc_dev",
    "qec_minimum_distance", "qec_minimum_distance_by_tests", "qec_sampler_sample_events24", "qec_sampler_sample_events16",
    "qec_dev_alloc", "qec_dev_free", "qec_dev_memset", "qec_dev_to_host", "qec_host_to_dev",
    "qec_host_alloc_pinned", "qec_host_free_pinned", "qec_stream_sync", "qec_last_kernel_ms",
    "qec_last_kernel", "qec_jit_compile_to_file", "qec_planner_signature_lists",
]

_lib = None


def tally_len(nc: int) -> int:
    return nc + 2 + TALLY_HIST_BINS


def lib() -> C.CDLL:
    """Load the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; "
                "g.build()'` (there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.qec_last_error.restype = C.c_char_p
        L.qec_version.restype = C.c_char_p
        L.qec_last_kernel.restype = C.c_char_p
        L.qec_circuit_bit_ops.restype = C.c_int64
        L.qec_circuit_bit_ops.argtypes = [C.c_void_p]
        L.qec_circuit_destroy.restype = None
        L.qec_circuit_destroy.argtypes = [C.c_void_p]
        L.qec_enumerator_destroy.restype = None
        L.qec_enumerator_destroy.argtypes = [C.c_void_p]
        L.qec_sampler_destroy.restype = None
        L.qec_sampler_destroy.argtypes = [C.c_void_p]
        _lib = L
    return _lib


def _check(rc: int) -> None:
    if rc == 0:
        return
    msg = lib().qec_last_error().decode()
    if rc == QEC_ERR_INVALID:
        raise ValueError(msg)
    raise RuntimeError(msg)


def _p(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _i32(x) -> np.ndarray:
    return np.ascontiguousarray(x, dtype=np.int32)


def device_count() -> int:
    return int(lib().qec_device_count())


def measure_int_peak(device: int = 0, ms: float = 50.0) -> Tuple[float, float]:
    """(INT32 logic lane-ops/s, nominal SM clock MHz) of ``device``."""
    ops, mhz = C.c_double(), C.c_double()
    _check(lib().qec_measure_int_peak(C.c_int(device), C.c_double(ms), C.byref(ops), C.byref(mhz)))
    return ops.value, mhz.value


def last_kernel_ms(device: int = 0) -> float:
    ms = C.c_float()
    _check(lib().qec_last_kernel_ms(C.c_int(device), C.byref(ms)))
    return ms.value


def last_kernel() -> str:
    """Name of the kernel that served this thread's latest enumerator / sampler launch
    ("...:specialised ..." = generated for the circuit, "...:table-driven (why)")."""
    return lib().qec_last_kernel().decode()


def jit_compile_to_file(nq: int, nc: int, encoded_circ, loc_op, kind: int, order: int = 1,
                        tally: bool = False, bits: bool = True, threads: int = 128,
                        min_blocks: int = 3, group: int = 8, cubin_path: str = "",
                        source_path: str = "") -> None:
    """Generate + compile (NVRTC, sm_100a) the specialised kernel of a circuit; no GPU needed."""
    off, data = rows_to_csr(encoded_circ)
    loc = _i32(loc_op)
    _check(lib().qec_jit_compile_to_file(
        C.c_int(nq), C.c_int(nc), C.c_int(len(encoded_circ)), _p(off), _p(data), C.c_int(len(loc)),
        _p(loc), C.c_int(kind), C.c_int(order), C.c_int(int(tally)), C.c_int(int(bits)),
        C.c_int(threads), C.c_int(min_blocks), C.c_int(group), cubin_path.encode(),
        source_path.encode()))


def planner_signature_lists(bits: np.ndarray, n_faults: int, pred=None, diff_basis: bool = False,
                            expect_short: bool = False):
    """The planners' host step on single-fault signatures ``bits[nc][ceil(n_faults / 32)]`` (no GPU
    needed): (list_off int64[n_faults + 1], entries uint16[...], raw_total, diff_total)."""
    bits = np.ascontiguousarray(bits, dtype=np.uint32)
    nc = bits.shape[0]
    assert bits.shape[1] == (n_faults + 31) // 32
    pr = None if pred is None else _i32(pred)
    off = np.zeros(n_faults + 1, dtype=np.int64)
    cap = max(1, int(nc) * int(n_faults))
    ent = np.zeros(cap, dtype=np.uint16)
    raw, dif = C.c_int64(), C.c_int64()
    _check(lib().qec_planner_signature_lists(_p(bits), C.c_int(nc), C.c_int64(n_faults), _p(pr),
                                             C.c_int(int(diff_basis)), C.c_int(int(expect_short)),
                                             _p(off), _p(ent), C.c_int64(cap), C.byref(raw), C.byref(dif)))
    return off, ent[:off[-1]], raw.value, dif.value


def rows_to_csr(encoded_circ: Sequence[Sequence[int]]) -> Tuple[np.ndarray, np.ndarray]:
    """The reference's ``encoded_circ`` (one list per op) as the C ABI takes it: int32 row offsets
    and the rows' entries back to back (``qec_circuit_create``)."""
    n = len(encoded_circ)
    off = np.zeros(n + 1, dtype=np.int32)
    if n:
        np.cumsum(np.fromiter(map(len, encoded_circ), dtype=np.int32, count=n), out=off[1:])
    data = np.fromiter(itertools.chain.from_iterable(encoded_circ), dtype=np.int32, count=int(off[-1]))
    return off, data


_PAULI_BITS = {"x": 1, "y": 3, "z": 2}


def pauli_code(pauli: str) -> int:
    """4-bit code of a Pauli string: char 0 -> operand a, char 1 -> operand b
    (reference errorpropagator.cpp:179-206; unknown characters and 'i' are identity)."""
    code = 0
    if len(pauli) > 0:
        code |= _PAULI_BITS.get(pauli[0], 0)
    if len(pauli) > 1:
        code |= _PAULI_BITS.get(pauli[1], 0) << 2
    return code


class Circuit:
    """``qec_circuit``: the compiled circuit IR resident on one GPU."""

    def __init__(self, nq: int, nc: int, encoded_circ: Sequence[Sequence[int]], device: int = 0):
        # encoded_circ: the reference's list of rows, or the (row_off, row_data) int32 arrays of
        # rows_to_csr -- the host buffers qec_circuit_create reads (a caller that creates many
        # objects of one circuit flattens the Python lists once)
        if isinstance(encoded_circ, tuple) and len(encoded_circ) == 2 and isinstance(encoded_circ[0], np.ndarray):
            off, data = _i32(encoded_circ[0]), _i32(encoded_circ[1])
            if len(off) < 1 or int(off[0]) != 0 or int(off[-1]) != len(data) or bool((off[1:] < off[:-1]).any()):
                raise ValueError("row offsets do not match the row data")
        else:
            off, data = rows_to_csr(encoded_circ)
        self.nq, self.nc, self.n_ops, self.device = int(nq), int(nc), len(off) - 1, int(device)
        h = C.c_void_p()
        _check(lib().qec_circuit_create(C.c_int(device), C.c_int(nq), C.c_int(nc),
                                        C.c_int(self.n_ops), _p(off), _p(data), C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            lib().qec_circuit_destroy(self._h)
            self._h = None

    __del__ = close

    @property
    def bit_ops(self) -> int:
        return int(lib().qec_circuit_bit_ops(self._h))

    def propagate_batch(self, path_off, fault_op, fault_pauli, want_frame: bool = False):
        """K1.  Returns packed outcomes uint32[nc][n_words] (and the final frame if asked)."""
        path_off = np.ascontiguousarray(path_off, dtype=np.int64)
        fault_op = _i32(fault_op)
        fault_pauli = np.ascontiguousarray(fault_pauli, dtype=np.uint8)
        n = len(path_off) - 1
        nw = (n + 31) // 32
        out = np.zeros((self.nc, nw), dtype=np.uint32)
        frame = np.zeros((2 * self.nq, nw), dtype=np.uint32) if want_frame else None
        _check(lib().qec_propagate_batch(self._h, C.c_int64(n), _p(path_off), _p(fault_op),
                                         _p(fault_pauli), _p(out), _p(frame)))
        return (out, frame) if want_frame else out


class Enumerator:
    """``qec_enumerator``."""

    def __init__(self, circuit: Circuit, order: int, loc_op, pauli_codes_per_loc):
        self.circuit, self.order = circuit, int(order)
        loc_op = _i32(loc_op)
        off = np.zeros(len(loc_op) + 1, dtype=np.int32)
        for i, codes in enumerate(pauli_codes_per_loc):
            off[i + 1] = off[i] + len(codes)
        flat = np.fromiter((c for codes in pauli_codes_per_loc for c in codes), dtype=np.uint8,
                           count=int(off[-1]))
        h = C.c_void_p()
        _check(lib().qec_enumerator_create(circuit._h, C.c_int(order), C.c_int(len(loc_op)),
                                           _p(loc_op), _p(off), _p(flat), C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            lib().qec_enumerator_destroy(self._h)
            self._h = None

    __del__ = close

    def num_paths(self) -> int:
        n = C.c_int64()
        _check(lib().qec_enumerator_num_paths(self._h, C.byref(n)))
        return n.value

    def unrank(self, index: int):
        locs = np.zeros(self.order, dtype=np.int32)
        digs = np.zeros(self.order, dtype=np.int32)
        _check(lib().qec_enumerator_unrank(self._h, C.c_int64(index), _p(locs), _p(digs)))
        return locs, digs

    def subset_rank(self, locs) -> Tuple[int, int]:
        locs = _i32(locs)
        first, count = C.c_int64(), C.c_int64()
        _check(lib().qec_enumerator_subset_rank(self._h, _p(locs), C.byref(first), C.byref(count)))
        return first.value, count.value

    def range(self, lo: int, hi: int, want_bits: bool = True, want_tally: bool = False):
        nw = (max(hi - lo, 0) + 31) // 32
        out = np.zeros((self.circuit.nc, nw), dtype=np.uint32) if want_bits else None
        tally = np.zeros(tally_len(self.circuit.nc), dtype=np.uint64) if want_tally else None
        _check(lib().qec_enumerator_range(self._h, C.c_int64(lo), C.c_int64(hi), _p(out), _p(tally)))
        return out, tally

    def range_dev(self, lo: int, hi: int, d_bits: int, d_tally: int, stream: int = 0):
        _check(lib().qec_enumerator_range_dev(self._h, C.c_int64(lo), C.c_int64(hi),
                                              C.c_void_p(d_bits or None), C.c_void_p(d_tally or None),
                                              C.c_void_p(stream or None)))


class Sampler:
    """``qec_sampler``."""

    def __init__(self, circuit: Circuit, loc_op, loc_class, class_p, class_codes, class_weights,
                 seed: int):
        self.circuit = circuit
        loc_op, loc_class = _i32(loc_op), _i32(loc_class)
        ncls = len(class_p)
        off = np.zeros(ncls + 1, dtype=np.int32)
        for i, codes in enumerate(class_codes):
            off[i + 1] = off[i] + len(codes)
        codes = np.fromiter((c for cs in class_codes for c in cs), dtype=np.uint8, count=int(off[-1]))
        w = np.fromiter((x for ws in class_weights for x in ws), dtype=np.float64, count=int(off[-1]))
        p = np.ascontiguousarray(class_p, dtype=np.float64)
        h = C.c_void_p()
        _check(lib().qec_sampler_create(circuit._h, C.c_int(len(loc_op)), _p(loc_op), _p(loc_class),
                                        C.c_int(ncls), _p(p), _p(off), _p(codes), _p(w),
                                        C.c_uint64(seed & (2**64 - 1)), C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            lib().qec_sampler_destroy(self._h)
            self._h = None

    __del__ = close

    def tables(self) -> np.ndarray:
        n = C.c_int()
        _check(lib().qec_sampler_tables(self._h, None, C.c_int(0), C.byref(n)))
        buf = np.zeros(n.value, dtype=np.uint32)
        _check(lib().qec_sampler_tables(self._h, _p(buf), C.c_int(n.value), C.byref(n)))
        return buf

    def sample(self, shot_lo: int, n_shots: int, want_bits: bool = True, want_tally: bool = False,
               fault_cap: int = 0):
        """Returns (bits, tally, fault_cnt, fault_ev); absent outputs are None."""
        nc = self.circuit.nc
        nw = (max(n_shots, 0) + 31) // 32
        out = np.zeros((nc, nw), dtype=np.uint32) if want_bits else None
        tally = np.zeros(tally_len(nc), dtype=np.uint64) if want_tally else None
        while True:
            cnt = np.zeros(nw, dtype=np.uint32) if fault_cap else None
            ev = np.zeros((nw, fault_cap), dtype=np.uint32) if fault_cap else None
            need = C.c_int()
            rc = lib().qec_sampler_sample(self._h, C.c_int64(shot_lo), C.c_int64(n_shots), _p(out),
                                          _p(tally), _p(cnt), _p(ev), C.c_int(fault_cap),
                                          C.byref(need))
            if rc == QEC_ERR_LIMIT and fault_cap and need.value > fault_cap:
                fault_cap = need.value + 8  # deterministic RNG: re-run with room for every event
                continue
            _check(rc)
            return out, tally, cnt, ev

    def sample_host_ptr(self, shot_lo: int, n_shots: int, out_ptr: int, tally_ptr: int = 0):
        """Host entry point with raw (ideally pinned) host addresses: kernel + D2H inside."""
        need = C.c_int()
        _check(lib().qec_sampler_sample(self._h, C.c_int64(shot_lo), C.c_int64(n_shots),
                                        C.c_void_p(out_ptr or None), C.c_void_p(tally_ptr or None),
                                        None, None, C.c_int(0), C.byref(need)))

    def sample_dev(self, shot_lo: int, n_shots: int, d_bits: int, d_tally: int, stream: int = 0):
        _check(lib().qec_sampler_sample_dev(self._h, C.c_int64(shot_lo), C.c_int64(n_shots),
                                            C.c_void_p(d_bits or None), C.c_void_p(d_tally or None),
                                            C.c_void_p(stream or None)))

    def check(self):
        _check(lib().qec_sampler_check(self._h))

    def sample_events(self, shot_lo: int, n_shots: int, want_tally: bool = False, pinned=None,
                      packed24: bool = False, packed16: bool = False):
        """Compact form: (word_cnt uint16[ceil(n/32)], events uint32[n_events], tally).  ``pinned``
        may hold preallocated (word_cnt, events) arrays (e.g. pinned host memory) to fill.
        ``packed24``: three bytes per event on the wire (events uint8[n_events, 3], see
        ``unpack_events24``); ``pinned`` then holds (word_cnt, uint8 buffer of 3 * capacity).
        ``packed16``: delta-coded 16-bit entries in slot order (``unpack_events16``; word_cnt then
        counts entries); ``pinned`` holds (word_cnt, uint16 buffer)."""
        nw = (max(n_shots, 0) + 31) // 32
        tally = np.zeros(tally_len(self.circuit.nc), dtype=np.uint64) if want_tally else None
        per = 3 if packed24 else 1                      # array elements per event / entry
        dt = np.uint8 if packed24 else np.uint16 if packed16 else np.uint32
        if pinned is not None:
            cnt, ev = pinned
        else:
            cnt = np.zeros(nw, dtype=np.uint16)
            ev = np.zeros(max(1024, int(1.2 * n_shots) + 4096) * per, dtype=dt)
        fn = (lib().qec_sampler_sample_events24 if packed24 else lib().qec_sampler_sample_events16 if packed16
              else lib().qec_sampler_sample_events)
        n_ev = C.c_int64()
        for _ in range(6):
            cap = len(ev) // per
            rc = fn(self._h, C.c_int64(shot_lo), C.c_int64(n_shots), _p(cnt), _p(ev), C.c_int64(cap),
                    C.byref(n_ev), _p(tally))
            if rc == QEC_ERR_OVERFLOW:
                continue                       # deterministic RNG: re-run with the raised capacity
            if rc == QEC_ERR_LIMIT and pinned is None and n_ev.value > cap:
                grow = int(n_ev.value * max(1.0, nw * 32 / max(1, n_shots)) * 2) + 4096
                ev = np.zeros(grow * per, dtype=dt)
                continue
            break
        _check(rc)
        if packed24:
            return cnt[:nw], ev[:3 * n_ev.value].reshape(-1, 3), tally
        return cnt[:nw], ev[:n_ev.value], tally

    def variant(self) -> str:
        """kernel variant and draw scheme (results are reproducible per variant)"""
        lib().qec_sampler_variant.restype = C.c_char_p
        return lib().qec_sampler_variant(self._h).decode()

    def counts(self, shot_lo: int, n_shots: int, table_cap: int = 0, want_tally: bool = False):
        """K4 on the hot path: (rows uint32[n][ceil(nc/32)], counts uint64[n], n_zero, tally) of
        the distinct NON-ZERO outcomes of the shots, histogrammed on the device; ``n_zero`` is
        the multiplicity of the all-zero outcome."""
        nc = self.circuit.nc
        rw = (nc + 31) // 32
        if table_cap <= 0:
            table_cap = 1 << 16
            while table_cap < 2 * min(n_shots, 1 << 26):
                table_cap <<= 1
        tally = np.zeros(tally_len(nc), dtype=np.uint64) if want_tally else None
        max_out = min(n_shots, table_cap)
        n_out, n_zero = C.c_int64(), C.c_uint64()
        # first pass sizes the output (distinct outcomes are usually far fewer than shots)
        probe = min(max_out, 1 << 16)
        while True:
            cnt = np.zeros(max(probe, 1), dtype=np.uint64)
            rows = np.zeros((max(probe, 1), rw), dtype=np.uint32)
            rc = lib().qec_sampler_counts(self._h, C.c_int64(shot_lo), C.c_int64(n_shots),
                                          C.c_int64(table_cap), C.c_int64(probe), _p(cnt), _p(rows),
                                          C.byref(n_out), C.byref(n_zero), _p(tally))
            if rc == QEC_ERR_LIMIT and n_out.value > probe and probe < max_out:
                probe = min(max_out, max(n_out.value, 2 * probe))  # deterministic RNG: re-run
                continue
            _check(rc)
            n = n_out.value
            return rows[:n], cnt[:n], int(n_zero.value), tally


class _Multi:
    """common part of the multi-GPU handles: one process, several devices (include/qec_b200.h)"""

    _destroy = None

    def close(self):
        if getattr(self, "_h", None):
            getattr(lib(), self._destroy)(self._h)
            self._h = None

    __del__ = close

    def _out(self, n, want_bits, want_tally):
        nw = (max(n, 0) + 31) // 32
        out = np.zeros((self.nc, nw), dtype=np.uint32) if want_bits else None
        tally = np.zeros(tally_len(self.nc), dtype=np.uint64) if want_tally else None
        return out, tally


class MultiSampler(_Multi):
    """``qec_multi_sampler``: FaultSampler on several GPUs of one box; results are one GPU's."""

    _destroy = "qec_multi_sampler_destroy"

    def __init__(self, devices, nq, nc, encoded_circ, loc_op, loc_class, class_p, class_codes, class_weights,
                 seed: int):
        self.nc, self.devices = int(nc), [int(d) for d in devices]
        lib().qec_multi_sampler_destroy.restype = None
        lib().qec_multi_sampler_destroy.argtypes = [C.c_void_p]
        roff, rdata = rows_to_csr(encoded_circ)
        loc_op, loc_class = _i32(loc_op), _i32(loc_class)
        off = np.zeros(len(class_p) + 1, dtype=np.int32)
        for i, codes in enumerate(class_codes):
            off[i + 1] = off[i] + len(codes)
        codes = np.fromiter((c for cs in class_codes for c in cs), dtype=np.uint8, count=int(off[-1]))
        w = np.fromiter((x for ws in class_weights for x in ws), dtype=np.float64, count=int(off[-1]))
        p = np.ascontiguousarray(class_p, dtype=np.float64)
        devs = _i32(self.devices)
        h = C.c_void_p()
        _check(lib().qec_multi_sampler_create(
            C.c_int(len(devs)), _p(devs), C.c_int(nq), C.c_int(nc), C.c_int(len(encoded_circ)), _p(roff),
            _p(rdata), C.c_int(len(loc_op)), _p(loc_op), _p(loc_class), C.c_int(len(class_p)), _p(p), _p(off),
            _p(codes), _p(w), C.c_uint64(seed & (2**64 - 1)), C.byref(h)))
        self._h = h

    def sample(self, shot_lo: int, n_shots: int, want_bits: bool = True, want_tally: bool = False):
        out, tally = self._out(n_shots, want_bits, want_tally)
        _check(lib().qec_multi_sampler_sample(self._h, C.c_int64(shot_lo), C.c_int64(n_shots), _p(out), _p(tally)))
        return out, tally

    def tally_dev(self, shot_lo: int, n_shots: int):
        """(tally, device ms): tallies only, asynchronous on every device + peer reduce"""
        _, tally = self._out(0, False, True)
        ms = C.c_float()
        _check(lib().qec_multi_sampler_tally_dev(self._h, C.c_int64(shot_lo), C.c_int64(n_shots), _p(tally),
                                                 C.byref(ms)))
        return tally, ms.value


class MultiEnumerator(_Multi):
    """``qec_multi_enumerator``: FaultEnumerator path ranges cut across several GPUs."""

    _destroy = "qec_multi_enumerator_destroy"

    def __init__(self, devices, nq, nc, encoded_circ, order, loc_op, pauli_codes_per_loc):
        self.nc, self.devices = int(nc), [int(d) for d in devices]
        lib().qec_multi_enumerator_destroy.restype = None
        lib().qec_multi_enumerator_destroy.argtypes = [C.c_void_p]
        roff, rdata = rows_to_csr(encoded_circ)
        loc_op = _i32(loc_op)
        off = np.zeros(len(loc_op) + 1, dtype=np.int32)
        for i, codes in enumerate(pauli_codes_per_loc):
            off[i + 1] = off[i] + len(codes)
        flat = np.fromiter((c for codes in pauli_codes_per_loc for c in codes), dtype=np.uint8, count=int(off[-1]))
        devs = _i32(self.devices)
        h = C.c_void_p()
        _check(lib().qec_multi_enumerator_create(
            C.c_int(len(devs)), _p(devs), C.c_int(nq), C.c_int(nc), C.c_int(len(encoded_circ)), _p(roff),
            _p(rdata), C.c_int(order), C.c_int(len(loc_op)), _p(loc_op), _p(off), _p(flat), C.byref(h)))
        self._h = h

    def num_paths(self) -> int:
        n = C.c_int64()
        _check(lib().qec_multi_enumerator_num_paths(self._h, C.byref(n)))
        return n.value

    def range(self, lo: int, hi: int, want_bits: bool = True, want_tally: bool = False):
        out, tally = self._out(hi - lo, want_bits, want_tally)
        _check(lib().qec_multi_enumerator_range(self._h, C.c_int64(lo), C.c_int64(hi), _p(out), _p(tally)))
        return out, tally

    def tally_dev(self, lo: int, hi: int):
        _, tally = self._out(0, False, True)
        ms = C.c_float()
        _check(lib().qec_multi_enumerator_tally_dev(self._h, C.c_int64(lo), C.c_int64(hi), _p(tally), C.byref(ms)))
        return tally, ms.value


def outcome_histogram_dev(circuit: Circuit, d_bits: int, n_paths: int, d_keys: int, d_counts: int,
                          table_cap: int, stream: int = 0, d_rows: int = 0, d_status: int = 0,
                          n_words: int = 0):
    """K4 over a device buffer (raw device addresses).  keys / counts / status are accumulated:
    zero them first."""
    _check(lib().qec_outcome_histogram_rows_dev(
        circuit._h, C.c_void_p(d_bits), C.c_int64(n_paths), C.c_int64(n_words or (n_paths + 31) // 32),
        C.c_void_p(d_keys), C.c_void_p(d_counts), C.c_void_p(d_rows or None), C.c_int64(table_cap),
        C.c_void_p(d_status or None), C.c_void_p(stream or None)))


def outcome_counts_noalloc_dev(circuit: Circuit, d_bits: int, n_paths: int, n_words: int, d_out_counts: int,
                               d_out_rows: int, max_out: int, d_out_status: int, stream: int = 0):
    """K4 with its hash table in the library's fixed device array: allocates nothing (raw device
    addresses in and out; see include/qec_b200.h for the limits and the sharing rule)."""
    _check(lib().qec_outcome_counts_noalloc_dev(
        circuit._h, C.c_void_p(d_bits), C.c_int64(n_paths), C.c_int64(n_words), C.c_void_p(d_out_counts or None),
        C.c_void_p(d_out_rows or None), C.c_int64(max_out), C.c_void_p(d_out_status), C.c_void_p(stream or None)))


def outcome_hash(rows: np.ndarray) -> np.ndarray:
    """The 64-bit key K4 gives an outcome (rows uint32[n][ceil(nc/32)]), for checking tables
    against numpy: mix64 folds of the 32-clbit tiles (csrc/qec_kernels.cu qec_mix64)."""
    h = np.full(len(rows), 0x243F6A8885A308D3, dtype=np.uint64)
    with np.errstate(over="ignore"):
        for j in range(rows.shape[1]):
            h ^= rows[:, j].astype(np.uint64)
            h *= np.uint64(0x9E3779B97F4A7C15)
            h ^= h >> np.uint64(29)
        h ^= h >> np.uint64(32)
    h[h == 0] = 1
    return h


def pack_rows(packed: np.ndarray, n: int) -> np.ndarray:
    """bit-sliced uint32[nc][n_words] -> row-packed uint32[n][ceil(nc/32)] (K4's outcome rows)"""
    nc = packed.shape[0]
    b = unpack_bits(packed, n)                                     # [n][nc]
    rw = (nc + 31) // 32
    pad = np.zeros((n, rw * 32), dtype=np.uint8)
    pad[:, :nc] = b
    return np.packbits(pad, axis=1, bitorder="little").view(np.uint32).reshape(n, rw)


def unpack_events24(ev24: np.ndarray) -> np.ndarray:
    """uint8[n, 3] events (location << 9 | shot bit << 4 | digit, little endian) -> the uint32 form
    location << 14 | shot bit << 9 | digit of ``sample_events``."""
    e = ev24[:, 0].astype(np.uint32) | (ev24[:, 1].astype(np.uint32) << 8) | (ev24[:, 2].astype(np.uint32) << 16)
    return ((e >> 9) << 14) | (((e >> 4) & 31) << 9) | (e & 15)


def unpack_events16(word_cnt: np.ndarray, entries: np.ndarray, tables: np.ndarray):
    """Delta-coded entries of ``sample_events(packed16=True)`` -> (event counts per word uint16,
    events uint32 in the form location << 14 | shot bit << 9 | digit, in slot order per word).
    ``tables`` = ``Sampler.tables()`` (carries the sorted -> original location map)."""
    B, K1, L = int(tables[0]), int(tables[1]), int(tables[3])
    slloc = tables[8 + K1 + B + L: 8 + K1 + B + 2 * L].astype(np.uint32)
    wc = word_cnt.astype(np.int64)
    e = entries.astype(np.int64)
    esc = (e >> 4) == 4095
    step = np.where(esc, 4095, (e >> 4) + 1)
    word_of = np.repeat(np.arange(len(wc)), wc)
    run = np.cumsum(step)
    first = np.concatenate([[0], np.cumsum(wc)])[:-1]           # index of each word's first entry
    base = np.where(wc > 0, run[np.minimum(first, len(run) - 1)] - step[np.minimum(first, len(run) - 1)], 0) if len(run) else first
    slot = run - base[word_of] - 1                               # slot of the entry inside its word
    keep = ~esc
    sl, bit = (slot[keep] >> 5), (slot[keep] & 31)
    events = (slloc[sl].astype(np.uint32) << 14) | (bit.astype(np.uint32) << 9) | (e[keep] & 15).astype(np.uint32)
    counts = np.bincount(word_of[keep], minlength=len(wc)).astype(np.uint16)
    return counts, events


def unpack_bits(packed: np.ndarray, n: int) -> np.ndarray:
    """bit-sliced uint32[nc][n_words] -> uint8[n][nc] (one row per path, as the reference's
    per-path outcome lists)."""
    nc, nw = packed.shape
    b = np.unpackbits(packed.view(np.uint8).reshape(nc, nw, 4), axis=2, bitorder="little")
    return np.ascontiguousarray(b.reshape(nc, nw * 32)[:, :n].T)
