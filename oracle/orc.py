"""TEST INFRASTRUCTURE ONLY — ctypes binding of the CPU oracle (oracle/pgq_oracle.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product package never does.  PARITY UNPINNED BY THE REFERENCE
(SURVEY.md §0): the oracle restates rules R1-R9 of SURVEY.md §8(a).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liborc.so")
_lib = None

I64P = C.POINTER(C.c_int64)
U8P = C.POINTER(C.c_uint8)


class Stats(C.Structure):
    _fields_ = [("batches", C.c_uint64), ("levels", C.c_uint64),
                ("edges_lane", C.c_uint64), ("edges_unique", C.c_uint64)]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


class ListEntry(C.Structure):
    _fields_ = [("offset", C.c_uint64), ("length", C.c_uint64)]


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".c", ".h"))]
    stale = (not os.path.exists(_SO)) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.orc_run_chunks_mt.restype = C.c_double
        _lib.orc_create_csr_vertex.restype = C.c_int64
        _lib.orc_shortestpath.restype = C.c_int
    return _lib


def _i64(a):
    a = np.ascontiguousarray(a, dtype=np.int64)
    return a, a.ctypes.data_as(I64P)


def _u8opt(a, n):
    if a is None:
        return None, None
    a = np.ascontiguousarray(a, dtype=np.uint8)
    assert a.shape[0] == n
    return a, a.ctypes.data_as(U8P)


def csr_build(V, src, dst, eid=None, want_eids=True):
    """-> (v[V+2], e[E], eids[E] or None), all int64, reference layout (R1-R3)."""
    src, ps = _i64(src)
    dst, pd = _i64(dst)
    E = src.shape[0]
    v = np.zeros(V + 2, dtype=np.int64)
    e = np.zeros(max(E, 1), dtype=np.int64)
    eids = np.zeros(max(E, 1), dtype=np.int64) if want_eids else None
    pe = None
    if eid is not None:
        eid, pe = _i64(eid)
    lib().orc_csr_build(C.c_int64(V), C.c_int64(E), ps, pd, pe, v.ctypes.data_as(I64P),
                        e.ctypes.data_as(I64P), eids.ctypes.data_as(I64P) if want_eids else None)
    return v, e[:E], (eids[:E] if want_eids else None)


def iterativelength(V, v, e, src, dst, src_valid=None, dst_valid=None, stats=False):
    v, pv = _i64(v)
    e, pe = _i64(e)
    src, ps = _i64(src)
    dst, pd = _i64(dst)
    n = src.shape[0]
    sv, psv = _u8opt(src_valid, n)
    dv, pdv = _u8opt(dst_valid, n)
    out = np.full(n, -1, dtype=np.int64)
    ov = np.zeros(n, dtype=np.uint8)
    st = Stats()
    lib().orc_iterativelength(C.c_int64(V), pv, pe, C.c_int64(n), ps, pd, psv, pdv,
                              out.ctypes.data_as(I64P), ov.ctypes.data_as(U8P),
                              C.byref(st) if stats else None)
    return (out, ov, st.as_dict()) if stats else (out, ov)


def iterativelength_bounded(V, v, e, src, dst, lower, upper, src_valid=None, dst_valid=None):
    """R11: at most `upper` levels (negative = unbounded); a shortest length below `lower` is NULL."""
    v, pv = _i64(v)
    e, pe = _i64(e)
    src, ps = _i64(src)
    dst, pd = _i64(dst)
    n = src.shape[0]
    sv, psv = _u8opt(src_valid, n)
    dv, pdv = _u8opt(dst_valid, n)
    out = np.full(n, -1, dtype=np.int64)
    ov = np.zeros(n, dtype=np.uint8)
    lib().orc_iterativelength_bounded(C.c_int64(V), pv, pe, C.c_int64(n), ps, pd, psv, pdv,
                                      C.c_int64(lower), C.c_int64(upper),
                                      out.ctypes.data_as(I64P), ov.ctypes.data_as(U8P), None)
    return out, ov


def reachability(V, v, e, src, dst, src_valid=None, dst_valid=None):
    v, pv = _i64(v)
    e, pe = _i64(e)
    src, ps = _i64(src)
    dst, pd = _i64(dst)
    n = src.shape[0]
    sv, psv = _u8opt(src_valid, n)
    dv, pdv = _u8opt(dst_valid, n)
    out = np.zeros(n, dtype=np.uint8)
    ov = np.zeros(n, dtype=np.uint8)
    lib().orc_reachability(C.c_int64(V), pv, pe, C.c_int64(n), ps, pd, psv, pdv,
                           out.ctypes.data_as(U8P), ov.ctypes.data_as(U8P), None)
    return out, ov


def shortestpath(V, v, e, eids, src, dst, src_valid=None, dst_valid=None, mode=0, faithful=False):
    """-> (offsets uint64[n], lengths uint64[n], valid uint8[n], values int64[total])."""
    v, pv = _i64(v)
    e, pe = _i64(e)
    peid = None
    if eids is not None:
        eids, peid = _i64(eids)
    src, ps = _i64(src)
    dst, pd = _i64(dst)
    n = src.shape[0]
    sv, psv = _u8opt(src_valid, n)
    dv, pdv = _u8opt(dst_valid, n)
    entries = (ListEntry * max(n, 1))()
    ov = np.zeros(n, dtype=np.uint8)
    vals = I64P()
    total = C.c_uint64(0)
    rc = lib().orc_shortestpath(C.c_int64(V), pv, pe, peid, C.c_int64(n), ps, pd, psv, pdv,
                                C.c_int(mode), C.c_int(1 if faithful else 0), entries,
                                ov.ctypes.data_as(U8P), C.byref(vals), C.byref(total), None)
    if rc != 0:
        raise RuntimeError("orc_shortestpath rc=%d" % rc)
    t = int(total.value)
    values = np.ctypeslib.as_array(vals, shape=(max(t, 1),))[:t].copy()
    lib().orc_free(vals)
    off = np.array([entries[i].offset for i in range(n)], dtype=np.uint64)
    ln = np.array([entries[i].length for i in range(n)], dtype=np.uint64)
    return off, ln, ov, values


def csr_weights(V, v, src, w):
    """Weights of the input rows in CSR slot order (R3 slot rule); dtype int64 or float64 kept."""
    v, pv = _i64(v)
    src, ps = _i64(src)
    w = np.ascontiguousarray(w)
    assert w.dtype in (np.int64, np.float64) and w.shape[0] == src.shape[0]
    out = np.empty_like(w)
    lib().orc_csr_weights(C.c_int64(V), C.c_int64(src.shape[0]), pv, ps, C.c_void_p(w.ctypes.data),
                          C.c_void_p(out.ctypes.data))
    return out


def cheapest_path_length(V, v, e, w, src, dst, src_valid=None, dst_valid=None):
    """R10 -> (out int64|float64 [n], valid uint8[n], rounds).  w in CSR slot order."""
    v, pv = _i64(v)
    e, pe = _i64(e)
    src, ps = _i64(src)
    dst, pd = _i64(dst)
    n = src.shape[0]
    sv, psv = _u8opt(src_valid, n)
    dv, pdv = _u8opt(dst_valid, n)
    w = np.ascontiguousarray(w)
    out = np.zeros(n, dtype=w.dtype)
    ov = np.zeros(n, dtype=np.uint8)
    rounds = C.c_uint64(0)
    fn = lib().orc_cheapest_path_length_i64 if w.dtype == np.int64 else lib().orc_cheapest_path_length_f64
    fn.restype = C.c_int
    rc = fn(C.c_int64(V), pv, pe, C.c_void_p(w.ctypes.data), C.c_int64(n), ps, pd, psv, pdv,
            C.c_void_p(out.ctypes.data), ov.ctypes.data_as(U8P), C.byref(rounds))
    if rc:
        raise ValueError("negative cycle")
    return out, ov, int(rounds.value)


def dijkstra(V, v, e, w, source):
    v, pv = _i64(v)
    e, pe = _i64(e)
    w, pw = _i64(w)
    dist = np.empty(V, dtype=np.int64)
    lib().orc_dijkstra_i64(C.c_int64(V), pv, pe, pw, C.c_int64(source), dist.ctypes.data_as(I64P))
    return dist


def bfs_simple(V, v, e, source):
    v, pv = _i64(v)
    e, pe = _i64(e)
    dist = np.zeros(max(V, 1), dtype=np.int32)
    lib().orc_bfs_simple(C.c_int64(V), pv, pe, C.c_int64(source), dist.ctypes.data_as(C.POINTER(C.c_int32)))
    return dist[:V]


def run_chunks_mt(kind, nthreads, V, v, e, src, dst, stats=False):
    """The reference's execution shape: <=1024-row calls on nthreads workers. -> (seconds, out, valid[, stats])"""
    v, pv = _i64(v)
    e, pe = _i64(e)
    src, ps = _i64(src)
    dst, pd = _i64(dst)
    n = src.shape[0]
    out = np.full(n, -1, dtype=np.int64) if kind == 0 else np.zeros(n, dtype=np.uint8)
    ov = np.zeros(n, dtype=np.uint8)
    st = Stats()
    secs = lib().orc_run_chunks_mt(C.c_int(kind), C.c_int(nthreads), C.c_int64(V), pv, pe, C.c_int64(n),
                                   ps, pd, out.ctypes.data_as(C.c_void_p), ov.ctypes.data_as(U8P),
                                   C.byref(st) if stats else None)
    return (secs, out, ov, st.as_dict()) if stats else (secs, out, ov)


def gen_rmat(seed, scale, E=None, row0=0, nrows=None, nthreads=8):
    E = (16 << scale) if E is None else E
    nrows = E if nrows is None else nrows
    src = np.empty(nrows, dtype=np.int64)
    dst = np.empty(nrows, dtype=np.int64)
    lib().orc_gen_rmat(C.c_uint64(seed), C.c_int(scale), C.c_int64(E), C.c_int64(row0), C.c_int64(nrows),
                       src.ctypes.data_as(I64P), dst.ctypes.data_as(I64P), C.c_int(nthreads))
    return src, dst


def gen_knows(seed, V, E, row0=0, nrows=None, nthreads=8):
    nrows = E if nrows is None else nrows
    src = np.empty(nrows, dtype=np.int64)
    dst = np.empty(nrows, dtype=np.int64)
    lib().orc_gen_knows(C.c_uint64(seed), C.c_int64(V), C.c_int64(E), C.c_int64(row0), C.c_int64(nrows),
                        src.ctypes.data_as(I64P), dst.ctypes.data_as(I64P), C.c_int(nthreads))
    return src, dst


def knows_params(V):
    m = C.c_uint64(0)
    a = C.c_uint64(0)
    lib().orc_knows_params(C.c_int64(V), C.byref(m), C.byref(a))
    return int(m.value), int(a.value)


def gen_pairs(seed, V, v, n):
    v, pv = _i64(v)
    src = np.empty(n, dtype=np.int64)
    dst = np.empty(n, dtype=np.int64)
    lib().orc_gen_pairs(C.c_uint64(seed), C.c_int64(V), pv, C.c_int64(n),
                        src.ctypes.data_as(I64P), dst.ctypes.data_as(I64P))
    return src, dst
