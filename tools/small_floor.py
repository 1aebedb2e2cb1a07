"""Dev helper: per-level floor of the small-graph kernel (chain graph: sweep work is negligible)."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pgq_b200
ctx = pgq_b200.Context(0)
for V in (64, 1024, 4736, 9472):
    src = np.arange(V - 1, dtype=np.int64)
    ctx.csr_build(1, V, src, src + 1)
    ctx.set_option("timing", 1)
    for rows in (1, 512):
        qs = np.zeros(rows, dtype=np.int64); qd = np.full(rows, 60, dtype=np.int64)
        ctx.iterativelength(1, qs, qd)
        ctx.reset_stats()
        for _ in range(10):
            ctx.iterativelength(1, qs, qd)
        st = ctx.stats()
        print("V %d rows %d: kernel %.1f us, levels %d -> %.2f us/level" % (V, rows, st["level_kernel_ms"] * 100, st["levels"] // 10, st["level_kernel_ms"] * 100 / (st["levels"] // 10)))
    ctx.csr_delete(1)
