"""Dev helper: build the scale-S R-MAT CSR (+ reverse CSR) a few times; for ncu launch lists."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pgq_b200
from bench import SEED_BASE
scale = int(sys.argv[1]) if len(sys.argv) > 1 else 22
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
ctx = pgq_b200.Context(0)
prep = True
for kv in sys.argv[3:]:  # option=value ... | noprep (forward build only)
    if kv == "noprep":
        prep = False
        continue
    k, val = kv.split("=")
    ctx.set_option(k, int(val))
V, E = 1 << scale, 16 << scale
d_src, d_dst = ctx.dev_alloc(E * 8), ctx.dev_alloc(E * 8)
ctx.gen_rmat_device(SEED_BASE + scale, scale, E, d_src, d_dst)
for _ in range(reps):
    ctx.csr_build_device(0, V, E, d_src, d_dst)
    print("csr_build_ms", ctx.stats()["csr_build_ms"])
    if prep:
        ctx.csr_prepare(0)
        print("transpose_ms", ctx.stats()["transpose_ms"])
    ctx.csr_delete(0)
