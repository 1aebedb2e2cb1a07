"""Dev helper: time the config-1 graph (LDBC-SF1 shape) call by call with the library's own timers."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pgq_b200
from bench import SEED_BASE, knows_params, make_pairs
V, E, n = 10000, 180000, 2048
args = [x for x in sys.argv[1:] if "=" in x]
for x in sys.argv[1:]:
    if x.startswith("V:"): V = int(x[2:])
    if x.startswith("E:"): E = int(x[2:])
ctx = pgq_b200.Context(0)
mult, add = knows_params(V)
d_src, d_dst = ctx.dev_alloc(E * 8), ctx.dev_alloc(E * 8)
ctx.gen_knows_device(SEED_BASE + 1, V, E, mult, add, d_src, d_dst)
src, dst = np.empty(E, dtype=np.int64), np.empty(E, dtype=np.int64)
ctx.d2h(src, d_src); ctx.d2h(dst, d_dst)
deg = np.bincount(src, minlength=V).astype(np.int64)
v_off = np.concatenate([[0], np.cumsum(deg), [E]]).astype(np.int64)
ps, pd = make_pairs(V, v_off, n, 77)
ctx.csr_build(1, V, src, dst)
for k, val in (kv.split('=') for kv in args):
    ctx.set_option(k, int(val))
ctx.iterativelength(1, ps[:1024], pd[:1024])
for rows in (1024, 512, 64, 1):
    t0 = time.perf_counter()
    for _ in range(50):
        ctx.iterativelength(1, ps[:rows], pd[:rows])
    print("rows %d: %.1f us per call (timers off)" % (rows, (time.perf_counter() - t0) / 50 * 1e6))
ctx.set_option("timing", 1); ctx.set_option("trace", 1)
for rows in (1024, 512, 64, 1):
    ctx.reset_stats()
    t0 = time.perf_counter()
    for _ in range(20):
        ctx.iterativelength(1, ps[:rows], pd[:rows])
    dt = (time.perf_counter() - t0) / 20
    st = ctx.stats()
    print("rows %d: %.1f us per call, kernel %.1f us per call, levels %d batches %d" %
          (rows, dt * 1e6, st["level_kernel_ms"] * 1e3 / 20, st["levels"] // 20, st["batches"] // 20))
    ctx.set_option("trace", 0)
