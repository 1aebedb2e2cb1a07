"""Dev helper: shortestpath on the config-4 shaped knows graph, a few batches (for ncu launch lists)."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pgq_b200
from bench import SEED_BASE, make_pairs, knows_params
V, E = 450000, 40000000
npairs = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
ctx = pgq_b200.Context(0)
mult, add = knows_params(V)
d_src, d_dst = ctx.dev_alloc(E * 8), ctx.dev_alloc(E * 8)
ctx.gen_knows_device(SEED_BASE + 100, V, E, mult, add, d_src, d_dst)
ctx.csr_build_device(0, V, E, d_src, d_dst, rowids=True)
ctx.csr_prepare(0)
v_off = np.empty(V + 2, dtype=np.int64)
ctx._chk(ctx.lib.pgq_csr_download(ctx._ctx, 0, v_off.ctypes.data_as(pgq_b200.I64P), None, None))
src, dst = make_pairs(V, v_off, npairs, 4242)
ctx.shortestpath(0, src, dst)
ctx.set_option("timing", 1)
if len(sys.argv) > 2:
    ctx.set_option("trace", 1)
ctx.reset_stats()
t0 = time.perf_counter()
r = ctx.shortestpath(0, src, dst)
print("shortestpath ms", (time.perf_counter() - t0) * 1e3, ctx.stats())
ctx.reset_stats()
t0 = time.perf_counter()
r = ctx.iterativelength(0, src, dst)
print("iterativelength ms", (time.perf_counter() - t0) * 1e3, ctx.stats())
