"""Dev helper: pairs/s of iterativelength at R-MAT scale S for different lane widths (words of W/8 bytes)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import pgq_b200
from bench import make_pairs, SEED_BASE
scale = int(sys.argv[1]) if len(sys.argv) > 1 else 24
npairs = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
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
ref = None
for lanes in (512, 256, 128, 64):
    for streams in (4, 8):
        ctx.set_option("lanes", lanes)
        ctx.set_option("streams", streams)
        out, ov = ctx.iterativelength(0, src, dst)
        if ref is None:
            ref = (out.copy(), ov.copy())
        assert np.array_equal(out, ref[0]) and np.array_equal(ov, ref[1])
        t0 = time.perf_counter()
        ctx.iterativelength(0, src, dst)
        dt = time.perf_counter() - t0
        print("lanes %3d streams %d: %8.1f ms  %9.0f pairs/s" % (lanes, streams, dt * 1e3, npairs / dt), flush=True)
