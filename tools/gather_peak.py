"""Dev helper: the memory system's rate for random 64-byte gathers (the dense level's access pattern),
for table sizes from L2-resident to far beyond L2.  Prints GB/s counting 68 bytes per gather."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pgq_b200
ctx = pgq_b200.Context(0)
G = 1 << 27  # 134 M gathers = 8.6 GB of words
for mb in (32, 128, 245, 512, 1024, 4096):
    nwords = mb * (1 << 20) // 64
    for u in (2, 4, 8):
        gbs, ms = ctx.probe_gather(nwords, G, u)
        print("table %5d MB  unroll %d  %8.1f GB/s  %7.3f ms" % (mb, u, gbs, ms), flush=True)
