"""Dev helper: per-level trace of one batch at a given scale (not part of the product)."""
import sys, numpy as np
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pgq_b200
from bench import make_pairs, SEED_BASE
scale = int(sys.argv[1]) if len(sys.argv) > 1 else 24
npairs = int(sys.argv[2]) if len(sys.argv) > 2 else 512
opts = dict(kv.split('=') for kv in sys.argv[3:])
ctx = pgq_b200.Context(0)
V, E = 1 << scale, 16 << scale
d_src, d_dst = ctx.dev_alloc(E * 8), ctx.dev_alloc(E * 8)
ctx.gen_rmat_device(SEED_BASE + scale, scale, E, d_src, d_dst)
ctx.csr_build_device(0, V, E, d_src, d_dst)
ctx.dev_free(d_src); ctx.dev_free(d_dst)
ctx.csr_prepare(0)
v_off = np.empty(V + 2, dtype=np.int64)
ctx._chk(ctx.lib.pgq_csr_download(ctx._ctx, 0, v_off.ctypes.data_as(pgq_b200.I64P), None, None))
src, dst = make_pairs(V, v_off, npairs, 1000)
for k, val in opts.items():
    ctx.set_option(k, int(val))
if os.environ.get("PGQ_TRACE_HOSTLOOP"):
    ctx.set_option("loop", 0)
ctx.iterativelength(0, src, dst)  # warm
ctx.set_option("timing", 1); ctx.set_option("trace", 1)
ctx.reset_stats()
ctx.iterativelength(0, src, dst)
print(ctx.stats())
