"""Dev helper: call-to-call jitter of shortestpath on the C1-shaped graph through the C ABI (options from argv)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import pgq_b200
from bench import SEED_BASE, knows_params, make_pairs
V, E, n = 10000, 180000, 2048
ctx = pgq_b200.Context(0)
mult, add = knows_params(V)
d_src, d_dst = ctx.dev_alloc(E * 8), ctx.dev_alloc(E * 8)
ctx.gen_knows_device(SEED_BASE + 1, V, E, mult, add, d_src, d_dst)
src, dst = np.empty(E, dtype=np.int64), np.empty(E, dtype=np.int64)
ctx.d2h(src, d_src); ctx.d2h(dst, d_dst)
deg = np.bincount(src, minlength=V).astype(np.int64)
v_off = np.concatenate([[0], np.cumsum(deg), [E]]).astype(np.int64)
ps, pd = make_pairs(V, v_off, n, 77)
for kv in sys.argv[1:]:
    k, v = kv.split("=")
    ctx.set_option(k, int(v))
for rep in range(3):
    ctx.csr_build(7, V, src, dst, rowids=True)
    for i in range(300):
        ctx.reachability(7, ps[i:i + 1], pd[i:i + 1])
    ts = []
    for k in range(12):
        t0 = time.perf_counter()
        for lo in range(0, n, 1024):
            ctx.shortestpath(7, ps[lo:lo + 1024], pd[lo:lo + 1024], mode=0)
        ts.append((time.perf_counter() - t0) * 1e3)
    tl = []
    for k in range(6):
        t0 = time.perf_counter()
        for lo in range(0, n, 1024):
            ctx.iterativelength(7, ps[lo:lo + 1024], pd[lo:lo + 1024])
        tl.append((time.perf_counter() - t0) * 1e3)
    print("rep %d %s shortestpath ms: %s | lengths ms: %s" % (rep, " ".join(sys.argv[1:]), " ".join("%.2f" % t for t in ts), " ".join("%.2f" % t for t in tl)), flush=True)
    ctx.csr_delete(7)
