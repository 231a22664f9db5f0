"""TEST INFRASTRUCTURE ONLY -- ctypes front end of the plain-C CPU restatement
(``oracle/qec_oracle.c``, ``oracle/sampler_port.c``).  Used as the checker in ``tests/``,
``__graft_entry__.smoke()`` and the "port" CPU baseline of ``bench.py``; never imported by
``qiskit_qec_b200``."""

from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "libqec_oracle.so")
_lib = None


def build() -> str:
    subprocess.run(["make", "-s", "-C", _HERE], check=True)
    return _LIB


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        L = C.CDLL(_LIB)
        L.orc_enumerate.restype = C.c_int64
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _csr(rows):
    off = np.zeros(len(rows) + 1, dtype=np.int32)
    for i, r in enumerate(rows):
        off[i + 1] = off[i] + len(r)
    data = np.array([v for r in rows for v in r], dtype=np.int32)
    return off, data


def _pstr(strings) -> np.ndarray:
    """list of Pauli strings -> NUL padded 3-byte records (only 2 chars are ever read)"""
    buf = np.zeros((max(len(strings), 1), 3), dtype=np.uint8)
    for i, s in enumerate(strings):
        b = s.encode()[:2]
        buf[i, :len(b)] = np.frombuffer(b, dtype=np.uint8)
    return buf


def propagate_batch(nq, nc, rows, paths, want_frame=False):
    """paths = [(icomb, [pauli strings])] -> uint8[n][nc] (and uint8[n][2nq])."""
    off, data = _csr(rows)
    poff = np.zeros(len(paths) + 1, dtype=np.int64)
    ops, strs = [], []
    for i, (icomb, errs) in enumerate(paths):
        poff[i + 1] = poff[i] + len(icomb)
        ops += list(icomb)
        strs += list(errs)
    fop = np.array(ops, dtype=np.int32)
    ps = _pstr(strs)
    out = np.zeros((len(paths), nc), dtype=np.uint8)
    frame = np.zeros((len(paths), 2 * nq), dtype=np.uint8) if want_frame else None
    rc = lib().orc_propagate_batch(C.c_int(nq), C.c_int(nc), C.c_int(len(rows)), _p(off), _p(data),
                                   C.c_int64(len(paths)), _p(poff), _p(fop), _p(ps), _p(out),
                                   _p(frame))
    if rc:
        raise RuntimeError("bad opcode")
    return (out, frame) if want_frame else out


def enumerate_range(nq, nc, rows, order, loc_op, paulis, lo=0, hi=None, want_paths=False):
    """All paths of the reference enumeration order with index in [lo, hi).
    Returns (total, outcomes uint8[hi-lo][nc][, locs, digits])."""
    off, data = _csr(rows)
    lop = np.array(loc_op, dtype=np.int32)
    po = np.zeros(len(paulis) + 1, dtype=np.int32)
    for i, pl in enumerate(paulis):
        po[i + 1] = po[i] + len(pl)
    ps = _pstr([p for pl in paulis for p in pl])
    if hi is None:
        hi = int(lib().orc_enumerate(C.c_int(nq), C.c_int(nc), C.c_int(len(rows)), _p(off),
                                     _p(data), C.c_int(order), C.c_int(len(lop)), _p(lop), _p(po),
                                     _p(ps), C.c_int64(0), C.c_int64(0), None, None, None))
        if hi < 0:
            raise ValueError("bad enumerator arguments")
    n = max(hi - lo, 0)
    out = np.zeros((n, nc), dtype=np.uint8)
    locs = np.zeros((n, order), dtype=np.int32) if want_paths else None
    digs = np.zeros((n, order), dtype=np.int32) if want_paths else None
    total = int(lib().orc_enumerate(C.c_int(nq), C.c_int(nc), C.c_int(len(rows)), _p(off), _p(data),
                                    C.c_int(order), C.c_int(len(lop)), _p(lop), _p(po), _p(ps),
                                    C.c_int64(lo), C.c_int64(hi), _p(out), _p(locs), _p(digs)))
    if total < 0:
        raise ValueError("bad enumerator arguments")
    return (total, out, locs, digs) if want_paths else (total, out)


def philox(counter, key):
    c = (C.c_uint32 * 4)(*counter)
    lib().orc_philox(c, C.c_uint32(key[0]), C.c_uint32(key[1]))
    return list(c)


_batch_cache = {}


def sample_batch(tables: np.ndarray, batch: int, cap: int = 1 << 16):
    """Accepted faults of one batch of the linear-response draw order: int array [n][3] of
    (sorted location, shot in batch, digit), in stream order."""
    ev = np.zeros(3 * cap, dtype=np.uint32)
    n = lib().orc_sample_batch(_p(tables), C.c_uint64(batch), _p(ev), C.c_int(cap))
    if n < 0:
        raise ValueError("tables are not those of a linear-response sampler")
    if n > cap:
        return sample_batch(tables, batch, n)
    return ev[:3 * n].reshape(n, 3).copy()


def sample_word(tables: np.ndarray, gw: int, nv: int = 32, cap: int = 4096):
    """Fault events (location << 14 | bit << 9 | digit) of global 32-shot word ``gw`` in slot
    order (sorted location, shot), for whichever draw order the tables describe."""
    if int(tables[7]) & 2:
        lgb = (int(tables[7]) >> 8) & 255
        batch, wb = gw >> (lgb - 5), gw & ((1 << (lgb - 5)) - 1)
        key = (tables.tobytes()[:64], int(tables.sum()), batch)
        if key not in _batch_cache:
            if len(_batch_cache) > 8:
                _batch_cache.clear()
            _batch_cache[key] = sample_batch(tables, batch)
        ev = _batch_cache[key]
        B, K1, L = int(tables[0]), int(tables[1]), int(tables[3])
        slloc = tables[8 + K1 + B + L: 8 + K1 + B + 2 * L]
        sel = ev[((ev[:, 1] >> 5) == wb) & ((ev[:, 1] & 31) < nv)]
        sel = sel[np.lexsort((sel[:, 1] & 31, sel[:, 0]))]
        return ((slloc[sel[:, 0]] << 14) | ((sel[:, 1] & 31) << 9) | sel[:, 2]).astype(np.uint32)
    ev = np.zeros(cap, dtype=np.uint32)
    n = lib().orc_sample_word(_p(tables), C.c_uint64(gw), C.c_int(nv), _p(ev), C.c_int(cap))
    if n > cap:
        return sample_word(tables, gw, nv, n)
    return ev[:n].copy()
