"""Dev helper: our CSR build against CUB's radix sort on the same rows (see cub_sort_ref.cu)."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import pgq_b200
from bench import SEED_BASE
lib = C.CDLL(os.path.join(os.path.dirname(os.path.abspath(__file__)), "libcubref.so"))
ctx = pgq_b200.Context(0)
for scale, bits in ((22, 22), (24, 24)):
    V, E = 1 << scale, 16 << scale
    d_src, d_dst = ctx.dev_alloc(E * 8), ctx.dev_alloc(E * 8)
    ctx.gen_rmat_device(SEED_BASE + scale, scale, E, d_src, d_dst)
    ctx.sync()
    s, nms = C.c_float(0), C.c_float(0)
    rc = lib.cubref_sort_ms(C.c_void_p(d_src), C.c_void_p(d_dst), C.c_longlong(E), C.c_int(bits), C.c_int(7), C.byref(s), C.byref(nms))
    ours = []
    for _ in range(5):
        ctx.csr_build_device(0, V, E, d_src, d_dst)
        ours.append(ctx.stats()["csr_build_ms"])
        ctx.csr_delete(0)
    print("scale %d: cub SortPairs(%d bits) %.3f ms (+ narrowing pass %.3f ms) rc=%d; our whole CSR build %.3f ms (median of 5: %s)" %
          (scale, bits, s.value, nms.value, rc, sorted(ours)[2], " ".join("%.2f" % x for x in ours)), flush=True)
    ctx.dev_free(d_src); ctx.dev_free(d_dst)
