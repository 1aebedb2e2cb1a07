"""Dev helper: cheapest_path_length on the C4-shaped knows graph, N pairs, options from argv (key=value).
   python tools/sssp_probe.py 64 sssp_slots=1"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import pgq_b200
from bench import SEED_BASE, knows_params, make_pairs
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
ctx = pgq_b200.Context(0)
V, E = 450000, 40000000
mult, add = knows_params(V)
d_src, d_dst = ctx.dev_alloc(E * 8), ctx.dev_alloc(E * 8)
ctx.gen_knows_device(SEED_BASE + 100, V, E, mult, add, d_src, d_dst)
ctx.csr_build_device(5, V, E, d_src, d_dst, rowids=True)
ctx.dev_free(d_src); ctx.dev_free(d_dst)
v_off = np.empty(V + 2, dtype=np.int64)
ctx._chk(ctx.lib.pgq_csr_download(ctx._ctx, 5, v_off.ctypes.data_as(pgq_b200.I64P), None, None))
src, dst = make_pairs(V, v_off, max(n, 1024), 7)
w = np.random.default_rng(4243).integers(1, 101, size=E).astype(np.int64)
if os.environ.get("PROBE_DOUBLE"):
    w = w.astype(np.float64) + 0.25  # fractional: the 8-byte double rows
ctx.csr_set_weights(5, w)
for kv in sys.argv[2:]:
    k, v = kv.split("=")
    ctx.set_option(k, int(v))
reps = 1 if os.environ.get("PROBE_ONCE") else 3
for r in range(reps):
    ctx.reset_stats()
    t0 = time.perf_counter()
    out, ov = ctx.cheapest_path_length(5, src[:n], dst[:n])
    ms = (time.perf_counter() - t0) * 1e3
    st = ctx.stats()
    print("pairs %d: %.2f ms, %.0f pairs/s, rounds %d batches %d narrow %d redo %d found %d" %
          (n, ms, n / ms * 1e3, st["levels"], st["batches"], st["sssp_narrow_batches"], st["sssp_redo_batches"],
           int(ov.sum())), flush=True)
