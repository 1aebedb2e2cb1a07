"""Multi-GPU host logic on CPU: source-batch sharding with world_size 2 over gloo.
The per-rank query is played by the oracle here (test infrastructure); on the GPU box the same
sharded_query() wraps Context.iterativelength (tests/test_gpu_parity.py::test_sharded_query_gpu)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_rows_partition():
    import pgq_b200
    for n in (0, 1, 511, 512, 513, 4096, 65536, 70001):
        for world in (1, 2, 3, 4, 8):
            spans = [pgq_b200.shard_rows(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for (a, b), (c, d) in zip(spans[:-1], spans[1:]):
                assert b == c and a <= b
            for a, b in spans:
                assert a % 512 == 0 or a == n
            sizes = [(b - a + 511) // 512 for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch.distributed as dist
    import orc
    import pgq_b200
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    scale = 11
    V = 1 << scale
    src, dst = orc.gen_rmat(0x5eed0000 + scale, scale)
    v, e, _ = orc.csr_build(V, src, dst)
    n = 2900
    ps, pd = orc.gen_pairs(3, V, v, n)

    def fn(s, d):
        return orc.iterativelength(V, v, e, s, d)

    out, valid = pgq_b200.sharded_query(fn, dist, n, world, rank, ps, pd)
    eo, ev = orc.iterativelength(V, v, e, ps, pd)
    ok = bool(np.array_equal(valid, ev) and np.array_equal(np.where(ev == 1, out, -1), np.where(ev == 1, eo, -1)))
    q.put((rank, ok))
    dist.destroy_process_group()


def test_sharded_query_world2_gloo():
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, True), (1, True)]


def test_part_ownership_is_a_partition():
    """1-D vertex partition: the scrambled block dealing covers every vertex once, each
    rank's local index space fits its slice of the replicated frontier structures, and R-MAT-style
    id-correlated weight ends up balanced."""
    import pgq_b200
    for V in (1, 31, 32, 33, 1000, 4096, 100003, 1 << 16):
        for world in (1, 2, 3, 8):
            vp = pgq_b200.part_slice(V, world)
            assert vp % 256 == 0 and vp * world >= V
            v = np.arange(V, dtype=np.int64)
            owner = pgq_b200.part_owner(v, V, world)
            local = pgq_b200.part_local(v, V, world)
            assert owner.min() >= 0 and owner.max() < world
            for r in range(world):
                mine = local[owner == r]
                assert len(np.unique(mine)) == len(mine)  # injective inside a rank
                assert len(mine) == 0 or mine.max() < vp
    # weight ~ 3^(number of zero bits): what R-MAT in-degree looks like
    V = 1 << 16
    v = np.arange(V, dtype=np.int64)
    zeros = 16 - np.array([bin(int(x)).count("1") for x in v])
    w = 3.0 ** zeros
    owner = pgq_b200.part_owner(v, V, 4)
    load = np.array([w[owner == r].sum() for r in range(4)])
    assert load.max() / load.min() < 1.6
