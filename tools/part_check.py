"""Partitioned (1-D, allgather-per-level) MS-BFS: parity + timing across ranks.

  torchrun --nproc-per-node G tools/part_check.py --scale 18 --pairs 600 [--lanes 64] [--check]

Every rank generates the same R-MAT edge stream on its GPU in chunks, keeps the rows whose dst it
owns (pgq_part_csr_add_edges_device), and all ranks answer the same pair set collectively.
--check compares with the CPU oracle on rank 0 (test infrastructure; small scales only)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pgq_b200  # noqa: E402
from bench import SEED_BASE, make_pairs  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=int, default=18)
    ap.add_argument("--edgefactor", type=int, default=16)
    ap.add_argument("--pairs", type=int, default=512)
    ap.add_argument("--lanes", type=int, default=64)
    ap.add_argument("--chunk", type=int, default=1 << 26)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--trace", action="store_true")
    ap.add_argument("--p2p", action="store_true", help="fused exchange over NVLink peer memory")
    ap.add_argument("--check-single", action="store_true",
                    help="rank 0 also answers the pairs with the (oracle-verified) single-GPU path and compares")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    uid = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        box = [pgq_b200.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        uid = box[0]
    ctx = pgq_b200.Context(local)
    ctx.part_init(rank, world, uid)
    V, E = 1 << a.scale, a.edgefactor << a.scale
    ctx.part_csr_create(0, V)
    chunk = min(a.chunk, E)
    d_src, d_dst = ctx.dev_alloc(chunk * 8), ctx.dev_alloc(chunk * 8)
    t0 = time.perf_counter()
    for row0 in range(0, E, chunk):
        n = min(chunk, E - row0)
        ctx._chk(ctx.lib.pgq_gen_rmat_rows_device(ctx._ctx, pgq_b200.C.c_uint64(SEED_BASE + a.scale),
                                                  pgq_b200.C.c_int(a.scale), pgq_b200.C.c_int64(row0),
                                                  pgq_b200.C.c_int64(n), pgq_b200.C.c_void_p(d_src),
                                                  pgq_b200.C.c_void_p(d_dst)))
        ctx.part_csr_add_edges_device(0, d_src, d_dst, n)
    ctx.part_csr_finalize(0)
    ctx.sync()
    build_s = time.perf_counter() - t0
    ctx.dev_free(d_src)
    ctx.dev_free(d_dst)
    # pairs: identical on every rank (needs out-degrees: use the oracle-free rule "any vertex")
    rng = np.random.RandomState(1234)
    ps = rng.randint(0, V, size=a.pairs).astype(np.int64)
    pd = rng.randint(0, V, size=a.pairs).astype(np.int64)
    ctx.set_option("lanes", a.lanes)
    if a.p2p and world > 1:
        ctx.part_p2p_setup(dist, 0, a.lanes, world)
    out, ov = ctx.part_iterativelength(0, ps, pd)  # warm
    ctx.set_option("timing", 1)
    if a.trace:
        ctx.set_option("trace", 1)
    ctx.reset_stats()
    ctx.sync()
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        out, ov = ctx.part_iterativelength(0, ps, pd)
    secs = (time.perf_counter() - t0) / a.steps
    st = ctx.stats()
    line = {"rank": rank, "world": world, "scale": a.scale, "V": V, "E": E, "pairs": a.pairs, "lanes": a.lanes,
            "ms_per_step": secs * 1e3, "pairs_per_s": a.pairs / secs, "levels": int(st["levels"] // a.steps),
            "batches": int(st["batches"] // a.steps), "pull_ms": st["level_kernel_ms"] / a.steps,
            "exchange_ms": st["coll_ms"] / a.steps, "coll_bytes_per_step": int(st["coll_bytes"] // a.steps),
            "build_s": build_s, "found": int(ov.sum()), "p2p": bool(a.p2p and world > 1)}
    if a.check and rank == 0:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import orc
        src, dst = orc.gen_rmat(SEED_BASE + a.scale, a.scale, E=E)
        v, e, _ = orc.csr_build(V, src, dst, want_eids=False)
        eo, ev = orc.iterativelength(V, v, e, ps, pd)
        ok = bool(np.array_equal(ov, ev) and np.array_equal(np.where(ev == 1, out, -1), np.where(ev == 1, eo, -1)))
        line["parity_vs_oracle"] = ok
    if a.check_single and rank == 0:
        d_s, d_d = ctx.dev_alloc(E * 8), ctx.dev_alloc(E * 8)
        ctx.gen_rmat_device(SEED_BASE + a.scale, a.scale, E, d_s, d_d)
        ctx.csr_build_device(1, V, E, d_s, d_d)
        ctx.dev_free(d_s)
        ctx.dev_free(d_d)
        ctx.set_option("lanes", 0)
        so, sv = ctx.iterativelength(1, ps, pd)
        line["parity_vs_single_gpu"] = bool(np.array_equal(sv, ov) and
                                            np.array_equal(np.where(sv == 1, so, -1), np.where(ov == 1, out, -1)))
    print(json.dumps(line), flush=True)
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()
    if a.check and rank == 0 and not line["parity_vs_oracle"]:
        sys.exit(1)


if __name__ == "__main__":
    main()
