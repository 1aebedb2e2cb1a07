"""Dev helper: where a config-1 sized CSR build (10 k vertices, 180 k edges, 1024-row chunks) spends its time."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import numpy as np
import pgq_b200, orc
V, E = 10000, 180000
src, dst = orc.gen_knows(7, V, E)
deg = np.bincount(src, minlength=V).astype(np.int64)
vid = np.arange(V, dtype=np.int64)
ctx = pgq_b200.Context(0)
for rep in range(4):
    t = [time.perf_counter()]
    ctx.create_csr_vertex(300, V, vid, deg); t.append(time.perf_counter())
    ctx.create_csr_edge(300, V, E, src[:1024], dst[:1024]); t.append(time.perf_counter())
    for lo in range(1024, E, 1024):
        ctx.create_csr_edge(300, V, E, src[lo:lo + 1024], dst[lo:lo + 1024])
    t.append(time.perf_counter())
    ctx.csr_finalize(300); t.append(time.perf_counter())
    ctx.iterativelength(300, src[:1024], dst[:1024]); t.append(time.perf_counter())
    ctx.iterativelength(300, src[:1024], dst[:1024]); t.append(time.perf_counter())
    ctx.csr_delete(300); t.append(time.perf_counter())
    names = ["vertex", "edge0", "edges", "finalize", "query1(+prepare)", "query2", "delete"]
    print("rep %d: " % rep + "  ".join("%s %.3f" % (n, (b - a) * 1e3) for n, a, b in zip(names, t, t[1:])), "ms; kernel csr_build_ms", ctx.stats()["csr_build_ms"], flush=True)
