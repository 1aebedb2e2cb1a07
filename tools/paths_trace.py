"""Dev helper: config-4 shaped graph (LDBC-SF100-like knows, V=450 k, E=40 M): per-level kernel times of one
512-lane batch and wall time of 16 384 pairs at 1/2/4/8 batches in flight, lengths and paths."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import pgq_b200
from bench import SEED_BASE, knows_params, make_pairs
V, E, npairs = 450000, 40000000, 16384
ctx = pgq_b200.Context(0)
mult, add = knows_params(V)
d_src, d_dst = ctx.dev_alloc(E * 8), ctx.dev_alloc(E * 8)
ctx.gen_knows_device(SEED_BASE + 100, V, E, mult, add, d_src, d_dst)
ctx.csr_build_device(200, V, E, d_src, d_dst, rowids=True)
ctx.dev_free(d_src); ctx.dev_free(d_dst)
ctx.csr_prepare(200)
v_off = np.empty(V + 2, dtype=np.int64)
ctx._chk(ctx.lib.pgq_csr_download(ctx._ctx, 200, v_off.ctypes.data_as(pgq_b200.I64P), None, None))
src, dst = make_pairs(V, v_off, npairs, 4242)
ctx.iterativelength(200, src[:512], dst[:512])
ctx.set_option("timing", 1); ctx.set_option("trace", 1)
ctx.reset_stats()
t0 = time.perf_counter(); ctx.iterativelength(200, src[:512], dst[:512]); t1 = time.perf_counter()
print("one batch, lengths: wall %.3f ms, level kernels %.3f ms" % ((t1 - t0) * 1e3, ctx.stats()["level_kernel_ms"]), flush=True)
ctx.reset_stats()
t0 = time.perf_counter(); ctx.shortestpath(200, src[:512], dst[:512]); t1 = time.perf_counter()
print("one batch, paths: wall %.3f ms, level kernels %.3f ms" % ((t1 - t0) * 1e3, ctx.stats()["level_kernel_ms"]), flush=True)
ctx.set_option("timing", 0); ctx.set_option("trace", 0)
for streams in (1, 2, 4, 8):
    ctx.set_option("streams", streams)
    for name, fn in (("lengths", lambda: ctx.iterativelength(200, src, dst)), ("paths", lambda: ctx.shortestpath(200, src, dst))):
        fn()
        t0 = time.perf_counter(); fn(); fn(); dt = (time.perf_counter() - t0) / 2
        print("streams %d %-8s %7.2f ms  %8.0f pairs/s" % (streams, name, dt * 1e3, npairs / dt), flush=True)
