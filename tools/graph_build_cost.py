"""Dev helper: what the device-side level loop costs a ONE-SHOT query -- fresh CSR, one 1024-row call, delete
(the DuckPGQ flow) -- against the host-driven loop, and what the first call pays for building the loop graph."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import pgq_b200
from bench import SEED_BASE, knows_params, make_pairs
ctx = pgq_b200.Context(0)
for V, E in ((450000, 40000000), (100000, 4000000), (1 << 20, 16 << 20)):
    mult, add = knows_params(V)
    d_src, d_dst = ctx.dev_alloc(E * 8), ctx.dev_alloc(E * 8)
    ctx.gen_knows_device(SEED_BASE + 100, V, E, mult, add, d_src, d_dst)
    for name, loop, minb in (("graph-always", 1, 0), ("host-loop", 0, 8), ("default", 1, 8)) * 2:
        ctx.set_option("loop", loop)
        ctx.set_option("loop_min_batches", minb)
        ctx.csr_build_device(5, V, E, d_src, d_dst)
        ctx.csr_prepare(5)
        v_off = np.empty(V + 2, dtype=np.int64)
        ctx._chk(ctx.lib.pgq_csr_download(ctx._ctx, 5, v_off.ctypes.data_as(pgq_b200.I64P), None, None))
        src, dst = make_pairs(V, v_off, 16384, 7)
        ts = []
        for k in range(7):
            t0 = time.perf_counter()
            ctx.iterativelength(5, src[k * 1024:(k + 1) * 1024], dst[k * 1024:(k + 1) * 1024])
            ts.append((time.perf_counter() - t0) * 1e3)
        t0 = time.perf_counter()
        ctx.iterativelength(5, src, dst)
        big = (time.perf_counter() - t0) * 1e3
        print("V=%d E=%d %s: 1024-row calls %s ms (first = fresh CSR), 16384-row call %.2f ms" %
              (V, E, name, " ".join("%.2f" % t for t in ts), big), flush=True)
        ctx.csr_delete(5)
    ctx.dev_free(d_src); ctx.dev_free(d_dst)
