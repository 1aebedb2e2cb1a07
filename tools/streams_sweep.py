"""Dev helper: iterativelength throughput against the `streams` option (batches in flight per call)."""
import sys, os, time, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pgq_b200
from bench import make_pairs, SEED_BASE, knows_params

ctx = pgq_b200.Context(0)


def graph(kind, cid):
    if kind == "rmat24":
        V, E = 1 << 24, 16 << 24
        d_src, d_dst = ctx.dev_alloc(E * 8), ctx.dev_alloc(E * 8)
        ctx.gen_rmat_device(SEED_BASE + 24, 24, E, d_src, d_dst)
    else:
        V, E = 450000, 40000000
        mult, add = knows_params(V)
        d_src, d_dst = ctx.dev_alloc(E * 8), ctx.dev_alloc(E * 8)
        ctx.gen_knows_device(SEED_BASE + 100, V, E, mult, add, d_src, d_dst)
    ctx.csr_build_device(cid, V, E, d_src, d_dst)
    ctx.dev_free(d_src); ctx.dev_free(d_dst)
    ctx.csr_prepare(cid)
    v_off = np.empty(V + 2, dtype=np.int64)
    ctx._chk(ctx.lib.pgq_csr_download(ctx._ctx, cid, v_off.ctypes.data_as(pgq_b200.I64P), None, None))
    return V, v_off


for kind, cid, npairs in (("rmat24", 0, 8192), ("knows", 1, 16384)):
    V, v_off = graph(kind, cid)
    src, dst = make_pairs(V, v_off, npairs, 1000)
    ref = None
    for streams in (1, 2, 3, 4):
        ctx.set_option("streams", streams)
        out = ctx.iterativelength(cid, src, dst)
        if ref is None:
            ref = out
        assert np.array_equal(out[0] * out[1], ref[0] * ref[1]) and np.array_equal(out[1], ref[1])
        ctx.sync()
        t0 = time.perf_counter()
        for _ in range(3):
            ctx.iterativelength(cid, src, dst)
        ms = (time.perf_counter() - t0) / 3 * 1e3
        print("%s pairs %d streams %d: %.2f ms  (%.0f pairs/s)" % (kind, npairs, streams, ms, npairs / ms * 1e3), flush=True)
    ctx.csr_delete(cid)
