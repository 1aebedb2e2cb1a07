"""pgq_b200 — Python host of the B200-native DuckPGQ graph hot path.

A thin ctypes caller of the C ABI in include/pgq_b200.h (libpgq_b200.so, hand-written CUDA
for sm_100a).  The method names and argument meaning mirror the reference's SQL scalar
functions: create_csr_vertex / create_csr_edge build a CSR keyed by an
integer csr_id; iterativelength / reachability / shortestpath answer batches of (src, dst)
rows with NULLs expressed as validity arrays.  There is NO CPU fallback: if the library or a
GPU is missing every compute call raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PGQ_B200_LIB") or os.path.join(_HERE, "libpgq_b200.so")  # (override: A/B builds)

I64P = C.POINTER(C.c_int64)
U8P = C.POINTER(C.c_uint8)

PGQ_CSR_ROWIDS = 1


class PgqError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("pgq_b200 error %d: %s" % (code, msg))
        self.code = code


class Stats(C.Structure):
    _fields_ = [("kernel_launches", C.c_uint64), ("batches", C.c_uint64), ("levels", C.c_uint64), ("push_levels", C.c_uint64),
                ("edges_lane", C.c_uint64), ("edges_unique", C.c_uint64), ("algo_bytes", C.c_uint64),
                ("level_kernel_ms", C.c_double), ("level_kernel_launches", C.c_uint64),
                ("csr_build_ms", C.c_double), ("transpose_ms", C.c_double),
                ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64),
                ("coll_bytes", C.c_uint64), ("coll_ms", C.c_double),
                ("model_bytes", C.c_uint64), ("k_listed", C.c_uint64), ("k_open", C.c_uint64), ("k_new", C.c_uint64),
                ("k_scanned", C.c_uint64), ("k_gathered", C.c_uint64), ("kp_items", C.c_uint64),
                ("kp_edges", C.c_uint64), ("kp_seen", C.c_uint64), ("kp_next", C.c_uint64),
                ("host_syncs", C.c_uint64), ("sssp_narrow_batches", C.c_uint64), ("sssp_redo_batches", C.c_uint64),
                ("sssp_narrow32_batches", C.c_uint64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class ListEntry(C.Structure):
    _fields_ = [("offset", C.c_uint64), ("length", C.c_uint64)]


_lib = None


def build(force=False):
    """Compile libpgq_b200.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    import subprocess
    args = ["make", "-C", _HERE, "-s", "-j8"]
    if force:
        args.append("-B")
    subprocess.check_call(args)
    return LIB_PATH


def load_library():
    """dlopen the CUDA library; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise PgqError(-2, "libpgq_b200.so not built (run __graft_entry__.build()); there is no CPU fallback")
        lib = C.CDLL(LIB_PATH)
        lib.pgq_last_error.restype = C.c_char_p
        lib.pgq_free.restype = None
        _lib = lib
    return _lib


def _i64(a):
    a = np.ascontiguousarray(a, dtype=np.int64)
    return a, a.ctypes.data_as(I64P)


def _u8(a, n):
    if a is None:
        return None, None
    a = np.ascontiguousarray(a, dtype=np.uint8)
    if a.shape[0] != n:
        raise ValueError("validity array length mismatch")
    return a, a.ctypes.data_as(U8P)


class Context:
    """One GPU context (one per process/rank).  Holds the csr_id -> CSR map like the reference's
    per-connection state."""

    def __init__(self, device=0, devices=None):
        """device: the GPU of a one-device context.  devices=[d0, d1, ...]: a multi-GPU context of ONE
        process (pgq_init_devices): CSRs are built on d0 and replicated on first use, every query deals
        its batches over all of them."""
        self.lib = load_library()
        self._ctx = C.c_void_p()
        if devices is not None:
            devices = [int(d) for d in devices]
            arr = (C.c_int * len(devices))(*devices)
            rc = self.lib.pgq_init_devices(arr, C.c_int(len(devices)), C.byref(self._ctx))
            device = devices[0] if devices else 0
        else:
            rc = self.lib.pgq_init(C.c_int(device), C.byref(self._ctx))
        if rc != 0:
            raise PgqError(rc, self.lib.pgq_last_error().decode())
        self.device = device
        self.devices = devices if devices is not None else [device]

    def _chk(self, rc):
        if rc != 0:
            raise PgqError(rc, self.lib.pgq_last_error().decode())

    def close(self):
        if self._ctx:
            self.lib.pgq_shutdown(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- options / stats ----
    def set_option(self, key, value):
        self._chk(self.lib.pgq_set_option(self._ctx, key.encode(), C.c_int64(int(value))))

    def stats(self):
        s = Stats()
        self._chk(self.lib.pgq_get_stats(self._ctx, C.byref(s)))
        return s.as_dict()

    def reset_stats(self):
        self._chk(self.lib.pgq_reset_stats(self._ctx))

    # ---- the reference's function surface ----
    def create_csr_vertex(self, csr_id, v_size, src_dense_id, cnt):
        """create_csr_vertex(csr_id, v_size, vertex id, out-degree) -> sum(cnt) (R2)."""
        if not getattr(self, "_vertex_started", {}).get(csr_id):
            self._chk(self.lib.pgq_csr_create(self._ctx, C.c_int32(csr_id), C.c_int64(v_size)))
            self.__dict__.setdefault("_vertex_started", {})[csr_id] = True
        vid, pv = _i64(src_dense_id)
        cnt, pc = _i64(cnt)
        out = C.c_int64(0)
        self._chk(self.lib.pgq_csr_set_degrees(self._ctx, C.c_int32(csr_id), pv, pc, C.c_int64(vid.shape[0]),
                                               C.byref(out)))
        return int(out.value)

    def create_csr_edge(self, csr_id, v_size, num_edges, src_dense_id, dst_dense_id, edge_rowid=None):
        """create_csr_edge(csr_id, v_size, num_edges, src, dst[, edge rowid]) -> rows accepted (R3).
        Rows are appended in call/arrival order, which is the stable key of the CSR."""
        st = self.__dict__.setdefault("_edge_started", {})
        if not st.get(csr_id):
            if not getattr(self, "_vertex_started", {}).get(csr_id):
                self._chk(self.lib.pgq_csr_create(self._ctx, C.c_int32(csr_id), C.c_int64(v_size)))
            self._chk(self.lib.pgq_csr_begin_edges(self._ctx, C.c_int32(csr_id), C.c_int64(num_edges)))
            st[csr_id] = True
        src, ps = _i64(src_dense_id)
        dst, pd = _i64(dst_dense_id)
        pe = None
        if edge_rowid is not None:
            eid, pe = _i64(edge_rowid)
        self._chk(self.lib.pgq_csr_add_edges(self._ctx, C.c_int32(csr_id), ps, pd, pe, C.c_int64(src.shape[0])))
        return int(src.shape[0])

    def _forget(self, csr_id):
        self.__dict__.get("_vertex_started", {}).pop(csr_id, None)
        self.__dict__.get("_edge_started", {}).pop(csr_id, None)

    def csr_finalize(self, csr_id):
        try:
            self._chk(self.lib.pgq_csr_finalize(self._ctx, C.c_int32(csr_id)))
        finally:
            self._forget(csr_id)

    def csr_build(self, csr_id, v_size, src, dst, eid=None, rowids=False):
        src, ps = _i64(src)
        dst, pd = _i64(dst)
        pe = None
        if eid is not None:
            eid, pe = _i64(eid)
        self._chk(self.lib.pgq_csr_build(self._ctx, C.c_int32(csr_id), C.c_int64(v_size), C.c_int64(src.shape[0]),
                                         ps, pd, pe, C.c_int(PGQ_CSR_ROWIDS if rowids else 0)))

    def csr_build_device(self, csr_id, v_size, e_size, d_src, d_dst, d_eid=0, rowids=False):
        self._chk(self.lib.pgq_csr_build_device(self._ctx, C.c_int32(csr_id), C.c_int64(v_size), C.c_int64(e_size),
                                                C.c_void_p(d_src), C.c_void_p(d_dst), C.c_void_p(d_eid or None),
                                                C.c_int(PGQ_CSR_ROWIDS if rowids else 0)))

    def csr_info(self, csr_id):
        v = C.c_int64(0)
        e = C.c_int64(0)
        self._chk(self.lib.pgq_csr_info(self._ctx, C.c_int32(csr_id), C.byref(v), C.byref(e)))
        return int(v.value), int(e.value)

    def csr_download(self, csr_id, want_eid=False):
        """-> (v[V+2], e[E], eid[E] or None) int64 arrays in the reference layout."""
        V, E = self.csr_info(csr_id)
        v = np.empty(V + 2, dtype=np.int64)
        e = np.empty(max(E, 1), dtype=np.int64)
        eid = np.empty(max(E, 1), dtype=np.int64) if want_eid else None
        self._chk(self.lib.pgq_csr_download(self._ctx, C.c_int32(csr_id), v.ctypes.data_as(I64P),
                                            e.ctypes.data_as(I64P),
                                            eid.ctypes.data_as(I64P) if want_eid else None))
        return v, e[:E], (eid[:E] if want_eid else None)

    def csr_delete(self, csr_id):
        self._forget(csr_id)
        self._chk(self.lib.pgq_csr_delete(self._ctx, C.c_int32(csr_id)))

    def csr_prepare(self, csr_id):
        self._chk(self.lib.pgq_csr_prepare(self._ctx, C.c_int32(csr_id)))

    def iterativelength(self, csr_id, src, dst, src_valid=None, dst_valid=None):
        """-> (length int64[n], valid uint8[n]); valid==0 is SQL NULL (R4/R5)."""
        src, ps = _i64(src)
        dst, pd = _i64(dst)
        n = src.shape[0]
        sv, psv = _u8(src_valid, n)
        dv, pdv = _u8(dst_valid, n)
        out = np.full(n, -1, dtype=np.int64)
        ov = np.zeros(n, dtype=np.uint8)
        self._chk(self.lib.pgq_iterativelength(self._ctx, C.c_int32(csr_id), C.c_int64(n), ps, pd, psv, pdv,
                                               out.ctypes.data_as(I64P), ov.ctypes.data_as(U8P)))
        return out, ov

    def iterativelength_bounded(self, csr_id, src, dst, lower, upper, src_valid=None, dst_valid=None):
        """6-argument iterativelength (R11): at most `upper` levels (negative = unbounded); a shortest
        length below `lower` is NULL."""
        src, ps = _i64(src)
        dst, pd = _i64(dst)
        n = src.shape[0]
        sv, psv = _u8(src_valid, n)
        dv, pdv = _u8(dst_valid, n)
        out = np.full(n, -1, dtype=np.int64)
        ov = np.zeros(n, dtype=np.uint8)
        self._chk(self.lib.pgq_iterativelength_bounded(self._ctx, C.c_int32(csr_id), C.c_int64(n), ps, pd, psv, pdv,
                                                       C.c_int64(lower), C.c_int64(upper),
                                                       out.ctypes.data_as(I64P), ov.ctypes.data_as(U8P)))
        return out, ov

    def csr_set_weights(self, csr_id, w, chunk=None):
        """Edge weights per INPUT ROW (int64 or float64), optionally sent in chunks like the rows of
        the 7-argument create_csr_edge would arrive (R10)."""
        w = np.ascontiguousarray(w)
        if w.dtype not in (np.int64, np.float64):
            raise ValueError("weights must be int64 or float64")
        wtype = 1 if w.dtype == np.int64 else 2
        n = w.shape[0]
        step = n if not chunk else int(chunk)
        for lo in range(0, max(n, 1), max(step, 1)):
            part = w[lo:lo + step]
            self._chk(self.lib.pgq_csr_set_weights(self._ctx, C.c_int32(csr_id), C.c_int(wtype),
                                                   C.c_void_p(part.ctypes.data), C.c_int64(lo),
                                                   C.c_int64(part.shape[0])))

    def csr_download_weights(self, csr_id):
        """Weights in CSR slot order (dtype by the weight type)."""
        wt = C.c_int(0)
        self._chk(self.lib.pgq_csr_weight_type(self._ctx, C.c_int32(csr_id), C.byref(wt)))
        if wt.value == 0:
            raise PgqError(-3, "csr has no weights")
        _, e_size = self.csr_info(csr_id)
        out = np.empty(e_size, dtype=np.int64 if wt.value == 1 else np.float64)
        self._chk(self.lib.pgq_csr_download_weights(self._ctx, C.c_int32(csr_id), C.c_void_p(out.ctypes.data)))
        return out

    def cheapest_path_length(self, csr_id, src, dst, src_valid=None, dst_valid=None):
        """-> (length int64|float64 [n], valid uint8[n]); NULL for NULL inputs / unreachable (R10)."""
        wt = C.c_int(0)
        self._chk(self.lib.pgq_csr_weight_type(self._ctx, C.c_int32(csr_id), C.byref(wt)))
        src, ps = _i64(src)
        dst, pd = _i64(dst)
        n = src.shape[0]
        sv, psv = _u8(src_valid, n)
        dv, pdv = _u8(dst_valid, n)
        out = np.zeros(n, dtype=np.float64 if wt.value == 2 else np.int64)
        ov = np.zeros(n, dtype=np.uint8)
        self._chk(self.lib.pgq_cheapest_path_length(self._ctx, C.c_int32(csr_id), C.c_int64(n), ps, pd, psv, pdv,
                                                    C.c_void_p(out.ctypes.data), ov.ctypes.data_as(U8P)))
        return out, ov

    def reachability(self, csr_id, src, dst, src_valid=None, dst_valid=None):
        src, ps = _i64(src)
        dst, pd = _i64(dst)
        n = src.shape[0]
        sv, psv = _u8(src_valid, n)
        dv, pdv = _u8(dst_valid, n)
        out = np.zeros(n, dtype=np.uint8)
        ov = np.zeros(n, dtype=np.uint8)
        self._chk(self.lib.pgq_reachability(self._ctx, C.c_int32(csr_id), C.c_int64(n), ps, pd, psv, pdv,
                                            out.ctypes.data_as(U8P), ov.ctypes.data_as(U8P)))
        return out, ov

    def shortestpath(self, csr_id, src, dst, src_valid=None, dst_valid=None, mode=0):
        """-> (offset uint64[n], length uint64[n], valid uint8[n], values int64[total]) (R7-R9)."""
        src, ps = _i64(src)
        dst, pd = _i64(dst)
        n = src.shape[0]
        sv, psv = _u8(src_valid, n)
        dv, pdv = _u8(dst_valid, n)
        entries = (ListEntry * max(n, 1))()
        ov = np.zeros(n, dtype=np.uint8)
        vals = I64P()
        total = C.c_uint64(0)
        self._chk(self.lib.pgq_shortestpath(self._ctx, C.c_int32(csr_id), C.c_int64(n), ps, pd, psv, pdv,
                                            C.c_int(mode), entries, ov.ctypes.data_as(U8P), C.byref(vals),
                                            C.byref(total)))
        t = int(total.value)
        values = np.ctypeslib.as_array(vals, shape=(max(t, 1),))[:t].copy() if vals else np.zeros(0, np.int64)
        if vals:
            self.lib.pgq_free(vals)
        ent = np.frombuffer(entries, dtype=np.uint64).reshape(-1, 2)[:n]
        return ent[:, 0].copy(), ent[:, 1].copy(), ov, values

    # ---- 1-D vertex partition (collective; include/pgq_b200.h) ----
    def part_init(self, rank, world, unique_id=None):
        self._chk(self.lib.pgq_part_init(self._ctx, C.c_int(rank), C.c_int(world),
                                         C.c_char_p(unique_id) if unique_id is not None else None))

    def part_range(self, v_size):
        base, cnt = C.c_int64(0), C.c_int64(0)
        self._chk(self.lib.pgq_part_range(self._ctx, C.c_int64(v_size), C.byref(base), C.byref(cnt)))
        return int(base.value), int(cnt.value)

    def part_csr_create(self, csr_id, v_size):
        self._chk(self.lib.pgq_part_csr_create(self._ctx, C.c_int32(csr_id), C.c_int64(v_size)))

    def part_csr_add_edges_device(self, csr_id, d_src, d_dst, n):
        self._chk(self.lib.pgq_part_csr_add_edges_device(self._ctx, C.c_int32(csr_id), C.c_void_p(d_src),
                                                         C.c_void_p(d_dst), C.c_int64(n)))

    def part_csr_finalize(self, csr_id):
        self._chk(self.lib.pgq_part_csr_finalize(self._ctx, C.c_int32(csr_id)))

    def part_p2p_setup(self, dist, csr_id, lanes, world):
        """Enable the fused exchange: export this rank's replica handles, all-gather them with the
        host plumbing (torch.distributed), map every peer's replicas (CUDA IPC over NVLink)."""
        self._chk(self.lib.pgq_part_p2p_close(self._ctx))  # mappings of an earlier setup go first, on every rank
        dist.barrier()
        mine = C.create_string_buffer(128)
        self._chk(self.lib.pgq_part_p2p_export(self._ctx, C.c_int32(csr_id), C.c_int(lanes), mine))
        box = [None] * world
        dist.all_gather_object(box, mine.raw)
        self._chk(self.lib.pgq_part_p2p_import(self._ctx, C.c_char_p(b"".join(box))))

    def part_iterativelength(self, csr_id, src, dst, src_valid=None, dst_valid=None):
        src, ps = _i64(src)
        dst, pd = _i64(dst)
        n = src.shape[0]
        sv, psv = _u8(src_valid, n)
        dv, pdv = _u8(dst_valid, n)
        out = np.empty(n, dtype=np.int64)
        ov = np.zeros(n, dtype=np.uint8)
        self._chk(self.lib.pgq_part_iterativelength(self._ctx, C.c_int32(csr_id), C.c_int64(n), ps, pd, psv, pdv,
                                                    out.ctypes.data_as(I64P), ov.ctypes.data_as(U8P)))
        return out, ov

    def part_reachability(self, csr_id, src, dst, src_valid=None, dst_valid=None):
        src, ps = _i64(src)
        dst, pd = _i64(dst)
        n = src.shape[0]
        sv, psv = _u8(src_valid, n)
        dv, pdv = _u8(dst_valid, n)
        out = np.zeros(n, dtype=np.uint8)
        ov = np.zeros(n, dtype=np.uint8)
        self._chk(self.lib.pgq_part_reachability(self._ctx, C.c_int32(csr_id), C.c_int64(n), ps, pd, psv, pdv,
                                                 out.ctypes.data_as(U8P), ov.ctypes.data_as(U8P)))
        return out, ov

    def iterativelength_device(self, csr_id, n, d_src, d_dst, d_out, d_out_valid):
        self._chk(self.lib.pgq_iterativelength_device(self._ctx, C.c_int32(csr_id), C.c_int64(n),
                                                      C.c_void_p(d_src), C.c_void_p(d_dst), C.c_void_p(d_out),
                                                      C.c_void_p(d_out_valid)))

    def iterativelength_device_noalloc(self, csr_id, n, d_src, d_dst, d_out, d_out_valid):
        """Allocation-free variant (static 8 MiB device scratch; small graphs, n <= 4096): include/pgq_b200.h."""
        self._chk(self.lib.pgq_iterativelength_device_noalloc(self._ctx, C.c_int32(csr_id), C.c_int64(n),
                                                              C.c_void_p(d_src), C.c_void_p(d_dst), C.c_void_p(d_out),
                                                              C.c_void_p(d_out_valid)))

    # ---- helpers ----
    def dev_alloc(self, nbytes):
        p = C.c_void_p()
        self._chk(self.lib.pgq_dev_alloc(self._ctx, C.c_uint64(nbytes), C.byref(p)))
        return p.value

    def dev_free(self, dptr):
        self._chk(self.lib.pgq_dev_free(self._ctx, C.c_void_p(dptr)))

    def h2d(self, dptr, arr):
        arr = np.ascontiguousarray(arr)
        self._chk(self.lib.pgq_dev_h2d(self._ctx, C.c_void_p(dptr), arr.ctypes.data_as(C.c_void_p),
                                       C.c_uint64(arr.nbytes)))

    def d2h(self, arr, dptr):
        assert arr.flags["C_CONTIGUOUS"]
        self._chk(self.lib.pgq_dev_d2h(self._ctx, arr.ctypes.data_as(C.c_void_p), C.c_void_p(dptr),
                                       C.c_uint64(arr.nbytes)))

    def sync(self):
        self._chk(self.lib.pgq_sync(self._ctx))

    def gen_rmat_device(self, seed, scale, e_size, d_src, d_dst):
        self._chk(self.lib.pgq_gen_rmat_device(self._ctx, C.c_uint64(seed), C.c_int(scale), C.c_int64(e_size),
                                               C.c_void_p(d_src), C.c_void_p(d_dst)))

    def probe_gather(self, nwords, ngathers, unroll=4):
        """-> (GB/s, ms) of random 64-byte gathers from a table of nwords words (measurement probe)."""
        g = C.c_double(0)
        ms = C.c_double(0)
        self._chk(self.lib.pgq_probe_gather(self._ctx, C.c_int64(nwords), C.c_int64(ngathers), C.c_int(unroll),
                                            C.byref(g), C.byref(ms)))
        return g.value, ms.value

    def gen_knows_device(self, seed, v_size, e_size, mult, add, d_src, d_dst):
        self._chk(self.lib.pgq_gen_knows_device(self._ctx, C.c_uint64(seed), C.c_int64(v_size), C.c_int64(e_size),
                                                C.c_uint64(mult), C.c_uint64(add), C.c_void_p(d_src),
                                                C.c_void_p(d_dst)))

    def flush_l2(self):
        self._chk(self.lib.pgq_flush_l2(self._ctx))

    def timer_start(self):
        self._chk(self.lib.pgq_timer_start(self._ctx))

    def timer_stop_ms(self):
        ms = C.c_double(0)
        self._chk(self.lib.pgq_timer_stop_ms(self._ctx, C.byref(ms)))
        return float(ms.value)


def pinned_empty(n, dtype):
    """numpy array over cudaHostAlloc'd memory (kept alive by the returned array's base)."""
    lib = load_library()
    dt = np.dtype(dtype)
    p = C.c_void_p()
    rc = lib.pgq_host_alloc_pinned(C.c_uint64(max(n, 1) * dt.itemsize), C.byref(p))
    if rc != 0:
        raise PgqError(rc, lib.pgq_last_error().decode())
    buf = (C.c_uint8 * (max(n, 1) * dt.itemsize)).from_address(p.value)
    arr = np.frombuffer(buf, dtype=dt)[:n]
    return arr, p.value


def pinned_free(hptr):
    load_library().pgq_host_free_pinned(C.c_void_p(hptr))


# ---- multi-GPU host logic: source-batch sharding ---------------------------------
def shard_rows(n, world, rank, lanes=512):
    """Rows [lo, hi) of an n-row pair set that `rank` answers.  Shards are contiguous runs of whole
    `lanes`-row batches (a batch is the independent unit: SURVEY.md §8(e)), sizes differing by at
    most one batch, so rank r's results are rows lo..hi of the caller's result vector."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad world/rank")
    batches = (n + lanes - 1) // lanes
    b_lo = batches * rank // world
    b_hi = batches * (rank + 1) // world
    return min(n, b_lo * lanes), min(n, b_hi * lanes)


def sharded_query(fn, dist, n, world, rank, src, dst, lanes=512):
    """Run `fn(src_shard, dst_shard) -> (out int64[], valid uint8[])` on this rank's shard and
    all-gather the shards into full-length arrays on every rank (torch.distributed: NCCL on GPUs,
    gloo in the CPU tests).  `fn` is the per-GPU query, e.g. Context.iterativelength bound to a
    csr id; nothing here touches the data path."""
    import torch
    lo, hi = shard_rows(n, world, rank, lanes)
    out, valid = fn(src[lo:hi], dst[lo:hi])
    if dist is None or world == 1:
        return np.asarray(out), np.asarray(valid)
    cap = max(shard_rows(n, world, r, lanes)[1] - shard_rows(n, world, r, lanes)[0] for r in range(world))
    buf = torch.full((cap, 2), -1, dtype=torch.int64)
    buf[:hi - lo, 0] = torch.from_numpy(np.ascontiguousarray(out, dtype=np.int64))
    buf[:hi - lo, 1] = torch.from_numpy(np.ascontiguousarray(valid, dtype=np.int64))
    dev = None
    if dist.get_backend() == "nccl":
        dev = torch.device("cuda", torch.cuda.current_device())
        buf = buf.to(dev)
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf)
    full_out = np.empty(n, dtype=np.int64)
    full_valid = np.empty(n, dtype=np.uint8)
    for r in range(world):
        a, b = shard_rows(n, world, r, lanes)
        p = parts[r].cpu().numpy()
        full_out[a:b] = p[:b - a, 0]
        full_valid[a:b] = p[:b - a, 1]
    return full_out, full_valid


def nccl_unique_id():
    """128-byte NCCL unique id (rank 0 creates it; broadcast it to the other ranks)."""
    buf = C.create_string_buffer(128)
    lib = load_library()
    rc = lib.pgq_nccl_unique_id(buf)
    if rc != 0:
        raise PgqError(rc, lib.pgq_last_error().decode())
    return buf.raw


def _part_kbits(v_size):
    nblk = (v_size + 31) // 32
    k = 0
    while (1 << k) < nblk:
        k += 1
    return k


def _part_mix(blk, k):
    """Bijection on k-bit block indices (host mirror of part_mix, csrc/common.cuh)."""
    blk = np.asarray(blk, dtype=np.uint64)
    if k < 2:
        return blk
    mask = np.uint64((1 << k) - 1)
    h = np.uint64((k + 1) // 2)
    with np.errstate(over="ignore"):
        x = blk ^ (blk >> h)
        x = (x * np.uint64(0x9E3779B97F4A7C15)) & mask
        x ^= x >> h
        x = (x * np.uint64(0xC2B2AE3D27D4EB4F)) & mask
        x ^= x >> h
    return x


def part_owner(v, v_size, world):
    """Owner rank of vertex v in the 1-D partition (scrambled 32-vertex blocks dealt round-robin)."""
    v = np.asarray(v, dtype=np.uint64)
    return (_part_mix(v >> np.uint64(5), _part_kbits(v_size)) % np.uint64(world)).astype(np.int64)


def part_local(v, v_size, world):
    v = np.asarray(v, dtype=np.uint64)
    m = _part_mix(v >> np.uint64(5), _part_kbits(v_size))
    return (((m // np.uint64(world)) << np.uint64(5)) | (v & np.uint64(31))).astype(np.int64)


def part_slice(v_size, world):
    """Vertices per rank in the replicated structures (multiple of 256)."""
    lblk = ((1 << _part_kbits(v_size)) + world - 1) // world
    return (lblk * 32 + 255) // 256 * 256
