"""pgq_iterativelength_device_noalloc (include/pgq_b200.h): the allocation-free variant, whose scratch is a static
8 MiB __device__ array, against the existing pgq_iterativelength_device, through the C ABI.  Collected last on
purpose.  Has its own context, so it passes alone in a fresh process as well as
after the other tests."""
import numpy as np
import pytest

import orc

pytestmark = pytest.mark.gpu


def _device_pairs(ctx, ps, pd):
    n = len(ps)
    bufs = [ctx.dev_alloc(n * 8), ctx.dev_alloc(n * 8), ctx.dev_alloc(n * 8), ctx.dev_alloc(n)]
    ctx.h2d(bufs[0], np.ascontiguousarray(ps, dtype=np.int64))
    ctx.h2d(bufs[1], np.ascontiguousarray(pd, dtype=np.int64))
    return bufs


def _run(ctx, fn, csr_id, bufs, n):
    # poison the outputs so stale results of an earlier call cannot pass for fresh ones
    ctx.h2d(bufs[2], np.full(n, -77, dtype=np.int64))
    ctx.h2d(bufs[3], np.full(n, 9, dtype=np.uint8))
    fn(csr_id, n, bufs[0], bufs[1], bufs[2], bufs[3])
    out = np.empty(n, dtype=np.int64)
    ov = np.empty(n, dtype=np.uint8)
    ctx.d2h(out, bufs[2])
    ctx.d2h(ov, bufs[3])
    return out, ov


def test_no_alloc_entry_point_matches_existing_one():
    import pgq_b200
    ctx = pgq_b200.Context(0)
    try:
        # graph 1: R-MAT scale 12 (4096 vertices): 512-lane words fit the static scratch
        scale = 12
        V = 1 << scale
        src, dst = orc.gen_rmat(0x5eed0000 + scale, scale)
        ctx.csr_build(41, V, src, dst)
        v, e, _ = orc.csr_build(V, src, dst, want_eids=False)
        with pytest.raises(pgq_b200.PgqError, match="pgq_csr_prepare"):
            b = _device_pairs(ctx, [1], [2])
            ctx.iterativelength_device_noalloc(41, 1, b[0], b[1], b[2], b[3])
        ctx.csr_prepare(41)
        # two calls in a row with different pair sets and lane widths (1500 rows = 3 batches of 512; 100 rows = 128
        # lanes): the second call sees the first one's leftovers in the shared scratch
        for seed, n in ((3, 1500), (4, 100), (5, 1500)):
            ps, pd = orc.gen_pairs(seed, V, v, n)
            ps[::37] = pd[::37]  # trivial rows
            bufs = _device_pairs(ctx, ps, pd)
            want = _run(ctx, ctx.iterativelength_device, 41, bufs, n)
            got = _run(ctx, ctx.iterativelength_device_noalloc, 41, bufs, n)
            again = _run(ctx, ctx.iterativelength_device_noalloc, 41, bufs, n)
            eo, ev = orc.iterativelength(V, v, e, ps, pd)
            for name, (o, val) in (("existing", want), ("noalloc", got), ("noalloc again", again)):
                assert np.array_equal(val, ev), name  # integer work: bit-exact, tolerance 0
                assert np.array_equal(np.where(ev == 1, o, -1), np.where(ev == 1, eo, -1)), name
            assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
            for p in bufs:
                ctx.dev_free(p)
        # graph 2: LDBC-SF1-shaped knows graph (config 1: 10 k vertices, 180 k edges), 1024-row call
        V2, E2 = 10000, 180000
        s2, d2 = orc.gen_knows(0x1dbc0001, V2, E2)
        ctx.csr_build(42, V2, s2, d2)
        ctx.csr_prepare(42)
        v2, e2, _ = orc.csr_build(V2, s2, d2, want_eids=False)
        ps, pd = orc.gen_pairs(1, V2, v2, 1024)
        bufs = _device_pairs(ctx, ps, pd)
        want = _run(ctx, ctx.iterativelength_device, 42, bufs, 1024)
        got = _run(ctx, ctx.iterativelength_device_noalloc, 42, bufs, 1024)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
        eo, ev = orc.iterativelength(V2, v2, e2, ps, pd)
        assert np.array_equal(got[1], ev) and np.array_equal(np.where(ev == 1, got[0], -1), np.where(ev == 1, eo, -1))
        # limits are errors that name themselves, never allocations
        with pytest.raises(pgq_b200.PgqError, match="limit is 4096"):
            ctx.iterativelength_device_noalloc(42, 5000, bufs[0], bufs[1], bufs[2], bufs[3])
        for p in bufs:
            ctx.dev_free(p)
        scale = 17  # 131072 vertices x 64 bytes x 3 = 24 MiB > 8 MiB
        src, dst = orc.gen_rmat(0x5eed0000 + scale, scale, E=4 << scale)
        ctx.csr_build(43, 1 << scale, src, dst)
        ctx.csr_prepare(43)
        ctx.set_option("lanes", 512)
        try:
            bufs = _device_pairs(ctx, [1, 2], [3, 4])
            with pytest.raises(pgq_b200.PgqError, match="8 MiB"):
                ctx.iterativelength_device_noalloc(43, 2, bufs[0], bufs[1], bufs[2], bufs[3])
            # out-of-range ids are reported like the existing entry point reports them
            ctx.h2d(bufs[0], np.array([1, 1 << 40], dtype=np.int64))
            ctx.set_option("lanes", 0)
            with pytest.raises(pgq_b200.PgqError, match="outside"):
                ctx.iterativelength_device_noalloc(41, 2, bufs[0], bufs[1], bufs[2], bufs[3])
        finally:
            ctx.set_option("lanes", 0)
    finally:
        ctx.close()
