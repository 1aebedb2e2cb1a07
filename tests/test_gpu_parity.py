"""GPU parity suite (-m gpu): the CUDA path, called through the C ABI, against the CPU oracle on
the same seeded inputs — bit-exact CSR arrays, lengths, reachability and paths."""
import json
import os

import numpy as np
import pytest

import orc
from golden_cases import CASES, CSR_EXPECT
from helpers import lengths_list, unpack_paths, assert_paths_equal, check_path_valid

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _build_both(ctx, V, src, dst, eid=None, csr_id=1):
    v, e, eids = orc.csr_build(V, src, dst, eid)
    ctx.csr_build(csr_id, V, src, dst, eid=eid, rowids=True)
    return v, e, eids


def _cmp_arr(name, a, b):
    a, b = np.asarray(a), np.asarray(b)
    if a.shape != b.shape or not np.array_equal(a, b):
        n = min(len(a), len(b))
        bad = np.nonzero(a[:n] != b[:n])[0]
        raise AssertionError("%s mismatch: shapes %s/%s first bad %s gpu=%s orc=%s" %
                             (name, a.shape, b.shape, bad[:8], a[bad[:8]], b[bad[:8]]))


@pytest.mark.parametrize("name", sorted(CASES))
def test_golden_cases(ctx, name):
    c = CASES[name]
    src = [a for a, _ in c["edges"]]
    dst = [b for _, b in c["edges"]]
    ctx.csr_build(7, c["V"], src, dst, rowids=True)
    v, e, eid = ctx.csr_download(7, want_eid=True)
    if name in CSR_EXPECT:
        assert v.tolist() == CSR_EXPECT[name]["v"]
        assert e.tolist() == CSR_EXPECT[name]["e"]
        assert eid.tolist() == CSR_EXPECT[name]["eid"]
    qs = [a for a, _ in c["q"]]
    qd = [b for _, b in c["q"]]
    out, ov = ctx.iterativelength(7, qs, qd)
    assert lengths_list(out, ov) == c["lengths"]
    r, rv = ctx.reachability(7, qs, qd)
    assert rv.tolist() == [1] * len(qs) and r.tolist() == c["reach"]
    assert unpack_paths(*ctx.shortestpath(7, qs, qd, mode=0)) == c["paths"]
    assert unpack_paths(*ctx.shortestpath(7, qs, qd, mode=1)) == c["paths_e"]
    ctx.csr_delete(7)


def test_cte_golden_upper_bound(ctx):
    """The reference DuckDB's own (recursive-CTE) answers with an upper bound on the path length."""
    g = json.load(open(os.path.join(GOLD, "cte_knows_bounded.json")))
    ctx.csr_build(8, g["V"], g["src"], g["dst"])
    top = str(max(g["uppers"]))
    out, ov = ctx.iterativelength(8, g["q_src"], g["q_dst"])
    assert lengths_list(out, ov) == g["expected"][top]
    r, rv = ctx.reachability(8, g["q_src"], g["q_dst"])
    assert r.tolist() == [int(x is not None) for x in g["expected"][top]]
    for u in g["uppers"]:
        out, ov = ctx.iterativelength_bounded(8, g["q_src"], g["q_dst"], 0, u)
        assert lengths_list(out, ov) == g["expected"][str(u)], u
    ctx.csr_delete(8)


def test_cte_golden(ctx):
    g = json.load(open(os.path.join(GOLD, "cte_lengths.json")))
    ctx.csr_build(8, g["V"], g["src"], g["dst"])
    out, ov = ctx.iterativelength(8, g["q_src"], g["q_dst"])
    assert lengths_list(out, ov) == g["expected"]
    ctx.csr_delete(8)


@pytest.mark.parametrize("scale,ef", [(8, 4), (12, 16), (16, 16), (18, 16)])
def test_csr_bit_identical_rmat(ctx, scale, ef):
    V = 1 << scale
    src, dst = orc.gen_rmat(0x5eed0000 + scale, scale, E=ef << scale)
    v, e, eids = _build_both(ctx, V, src, dst)
    gv, ge, geid = ctx.csr_download(1, want_eid=True)
    _cmp_arr("v", gv, v)
    _cmp_arr("e", ge, e)
    _cmp_arr("eid", geid, eids)
    ctx.csr_delete(1)


@pytest.mark.parametrize("wide,lookback", [(1, 1), (1, 0), (2, 1), (0, 0), (3, 1), (3, 0)])
@pytest.mark.parametrize("V,E", [(2000, 40000), ((1 << 11) + 1, 50000), ((1 << 17) - 5, 300000),
                                 ((1 << 22) - 3, 250000), ((1 << 22) + 9, 100000), (3, 70000),
                                 ((1 << 23) + 1, 150000), ((1 << 24) - 7, 200000), ((1 << 24) + 5, 120000)])
def test_csr_both_sort_cores(ctx, V, E, wide, lookback):
    """The two-pass 11-bit sort (keys <= 22 bits; both block shapes; tile bases by decoupled look-back or by
    histogram pre-pass + scan) and the 8-bit sort give the same, oracle-identical CSR:
    digit splits 6+5 / 6+6 / 9+8 / 11+11, one pass (<= 11 bits), 23 / 24 bits (sort_wide=3: two 12-bit passes, 4096
    buckets, 512-thread blocks; otherwise the 8-bit core), > 24 bits (always the 8-bit core), ragged
    last tiles, heavy duplicates (V = 3), forward arrays + edge ids and the reverse CSR (checked through
    a query)."""
    rng = np.random.default_rng(V + E)
    src = rng.integers(0, V, E, dtype=np.int64)
    dst = rng.integers(0, V, E, dtype=np.int64)
    hot = rng.integers(0, V, 8)
    src[::5] = hot[rng.integers(0, 8, len(src[::5]))]  # a few very long neighbour lists
    eid = rng.permutation(E).astype(np.int64) * 3 + 1
    v, e, eids = orc.csr_build(V, src, dst, eid)
    ctx.set_option("sort_wide", wide)
    ctx.set_option("sort_lookback", lookback)
    ctx.set_option("sort_wide_min", 0)
    try:
        ctx.csr_build(2, V, src, dst, eid=eid)
        gv, ge, geid = ctx.csr_download(2, want_eid=True)
        _cmp_arr("v", gv, v)
        _cmp_arr("e", ge, e)
        _cmp_arr("eid", geid, eids)
        big = V > (1 << 20)  # V x 64-byte bitmaps per level: keep the query small there
        ps, pd = orc.gen_pairs(3, V, v, 40 if big else 300)
        got = ctx.iterativelength(2, ps, pd) if big else ctx.shortestpath(2, ps, pd, mode=1)
    finally:
        ctx.set_option("sort_wide", 1)
        ctx.set_option("sort_lookback", 1)
        ctx.set_option("sort_wide_min", 32768)
    if big:
        eo, ev = orc.iterativelength(V, v, e, ps, pd)
        _cmp_arr("valid", got[1], ev)
        _cmp_arr("length", got[0], eo)
    else:
        assert_paths_equal(got, orc.shortestpath(V, v, e, eids, ps, pd, mode=1), "paths after sort core %d" % wide)
    ctx.csr_delete(2)


def test_small_block_cache_reuse(ctx):
    """CSRs built, queried and deleted back to back run on recycled device / pinned blocks (api.cu small-block
    cache): recycled memory is not zero like a fresh cudaMalloc often is, so every array must be fully written."""
    rng = np.random.default_rng(4)
    for rep in range(6):
        V = int(rng.integers(50, 6000))
        E = int(rng.integers(1, 40000))
        src = rng.integers(0, V, E, dtype=np.int64)
        dst = rng.integers(0, V, E, dtype=np.int64)
        v, e, eids = orc.csr_build(V, src, dst)
        deg = np.bincount(src, minlength=V).astype(np.int64)
        ctx.create_csr_vertex(9, V, np.arange(V), deg)
        for lo in range(0, E, 1024):
            ctx.create_csr_edge(9, V, E, src[lo:lo + 1024], dst[lo:lo + 1024])
        ctx.csr_finalize(9)
        gv, ge, _ = ctx.csr_download(9)
        _cmp_arr("v", gv, v)
        _cmp_arr("e", ge, e)
        ps, pd = orc.gen_pairs(rep, V, v, 700)
        out, ov = ctx.iterativelength(9, ps, pd)
        eo, ev = orc.iterativelength(V, v, e, ps, pd)
        _cmp_arr("valid", ov, ev)
        _cmp_arr("length", out, eo)
        assert_paths_equal(ctx.shortestpath(9, ps[:200], pd[:200], mode=0),
                           orc.shortestpath(V, v, e, eids, ps[:200], pd[:200], mode=0), "paths rep %d" % rep)
        ctx.csr_delete(9)
        if rep == 3:
            ctx.set_option("trim_cache", 1)


def test_csr_with_explicit_edge_ids_and_chunked_rows(ctx):
    """create_csr_vertex + create_csr_edge fed in 1024-row chunks like DuckDB does (R2/R3)."""
    V, E = 5000, 60000
    src, dst = orc.gen_knows(21, V, E)
    eid = (np.arange(E, dtype=np.int64) * 7 + 13)
    v, e, eids = orc.csr_build(V, src, dst, eid)
    cnt = np.bincount(src, minlength=V).astype(np.int64)
    total = 0
    for lo in range(0, V, 1024):
        total += ctx.create_csr_vertex(3, V, np.arange(lo, min(V, lo + 1024)), cnt[lo:lo + 1024])
    assert total == E
    for lo in range(0, E, 1024):
        ctx.create_csr_edge(3, V, E, src[lo:lo + 1024], dst[lo:lo + 1024], eid[lo:lo + 1024])
    ctx.csr_finalize(3)
    gv, ge, geid = ctx.csr_download(3, want_eid=True)
    _cmp_arr("v", gv, v)
    _cmp_arr("e", ge, e)
    _cmp_arr("eid", geid, eids)
    ctx.csr_delete(3)


def test_csr_count_mismatch_is_an_error(ctx):
    import pgq_b200
    ctx.create_csr_vertex(4, 4, [0, 1, 2, 3], [1, 0, 0, 0])
    ctx.create_csr_edge(4, 4, 2, [0, 1], [1, 2])
    with pytest.raises(pgq_b200.PgqError):
        ctx.csr_finalize(4)
    ctx.csr_delete(4)
    with pytest.raises(pgq_b200.PgqError):
        ctx.csr_build(4, 4, [0, 9], [1, 2])  # id out of range
    ctx.csr_delete(4)


def test_empty_graphs(ctx):
    ctx.csr_build(5, 3, [], [])
    v, e, _ = ctx.csr_download(5)
    assert v.tolist() == [0, 0, 0, 0, 0] and len(e) == 0
    out, ov = ctx.iterativelength(5, [0, 1], [1, 1])
    assert lengths_list(out, ov) == [None, 0]
    out, ov = ctx.iterativelength(5, [], [])
    assert len(out) == 0
    ctx.csr_delete(5)


@pytest.mark.parametrize("lanes", [0, 64, 128, 256, 512])
def test_lengths_rmat_all_lane_widths(ctx, lanes):
    scale = 14
    V = 1 << scale
    src, dst = orc.gen_rmat(0x5eed0000 + scale, scale)
    v, e, _ = _build_both(ctx, V, src, dst)
    n = 1500
    ps, pd = orc.gen_pairs(5, V, v, n)
    rng = np.random.default_rng(1)
    sv = (rng.random(n) > 0.01).astype(np.uint8)
    dv = (rng.random(n) > 0.01).astype(np.uint8)
    ps[::97] = pd[::97]  # ~1% trivial rows
    ctx.set_option("lanes", lanes)
    ctx.set_option("stats", 1)
    ctx.reset_stats()
    try:
        out, ov = ctx.iterativelength(1, ps, pd, sv, dv)
        st = ctx.stats()
    finally:
        ctx.set_option("lanes", 0)
        ctx.set_option("stats", 0)
    eo, ev, est = orc.iterativelength(V, v, e, ps, pd, sv, dv, stats=True)
    _cmp_arr("valid", ov, ev)
    _cmp_arr("length", out, eo)
    if lanes in (0, 512):
        # TEPS numerator and batch/level structure agree with the reference-shaped oracle
        assert st["edges_lane"] == est["edges_lane"]
        assert st["batches"] == est["batches"] and st["levels"] <= est["levels"]
        assert st["edges_unique"] == est["edges_unique"]
    r, rv = ctx.reachability(1, ps, pd, sv, dv)
    er, erv = orc.reachability(V, v, e, ps, pd, sv, dv)
    _cmp_arr("reach valid", rv, erv)
    _cmp_arr("reach", r, er)
    ctx.csr_delete(1)


@pytest.mark.parametrize("graph", ["rmat13", "knows", "chain"])
def test_bounded_lengths(ctx, graph):
    """6-argument iterativelength (R11); the ctx fixture runs it on the whole-batch kernel and on the
    general level loop."""
    if graph == "rmat13":
        V = 1 << 13
        src, dst = orc.gen_rmat(0x5eed0000 + 13, 13)
    elif graph == "knows":
        V = 3000
        src, dst = orc.gen_knows(31, V, 30000)
    else:  # long levels: the cap has to stop the loop mid-way
        V = 200
        src, dst = np.arange(V - 1), np.arange(1, V)
    v, e, _ = _build_both(ctx, V, src, dst)
    n = 1400
    ps, pd = orc.gen_pairs(17, V, v, n)
    if graph == "chain":
        ps, pd = np.minimum(ps, pd), np.maximum(ps, pd)
    ps[::61] = pd[::61]
    sv = (np.arange(n) % 37 != 0).astype(np.uint8)
    for lo, hi in [(0, 0), (0, 1), (0, 2), (1, 3), (2, 4), (3, 3), (0, 60), (4, 1), (0, -1)]:
        out, ov = ctx.iterativelength_bounded(1, ps, pd, lo, hi, sv)
        eo, ev = orc.iterativelength_bounded(V, v, e, ps, pd, lo, hi, sv)
        _cmp_arr("valid %d..%d" % (lo, hi), ov, ev)
        _cmp_arr("length %d..%d" % (lo, hi), out, eo)
    # the cap is per call: the unbounded function afterwards sees the whole search again
    out, ov = ctx.iterativelength(1, ps, pd, sv)
    eo, ev = orc.iterativelength(V, v, e, ps, pd, sv)
    _cmp_arr("valid after", ov, ev)
    _cmp_arr("length after", out, eo)
    ctx.csr_delete(1)


@pytest.mark.parametrize("bins", [(0, 0, 0, 1), (1, 2, 4, 8), (4, 32, 64, 4096), (0, 16, 32, 64),
                                  (1 << 30, 1 << 30, 1 << 30, 1 << 30), (2, 1 << 30, 1 << 30, 1 << 30),
                                  (2, 4, 1 << 30, 1 << 30)])
@pytest.mark.parametrize("lanes", [64, 256, 512])
def test_lengths_all_degree_bins(ctx, bins, lanes):
    """Force every vertex through each in-degree class kernel (tiny / small / medium / warp / hub), with and
    without the frontier-bitmap filter, at every threads-per-word layout."""
    V, E = 3000, 90000
    src, dst = orc.gen_knows(33, V, E)
    v, e, _ = _build_both(ctx, V, src, dst)
    ps, pd = orc.gen_pairs(6, V, v, 700)
    eo, ev = orc.iterativelength(V, v, e, ps, pd)
    _, _, eids = orc.csr_build(V, src, dst)
    ep = orc.shortestpath(V, v, e, eids, ps, pd, mode=1)
    ctx.set_option("tiny_max", bins[0])
    ctx.set_option("small_max", bins[1])
    ctx.set_option("med_max", bins[2])
    ctx.set_option("hub_min", bins[3])
    ctx.set_option("hub_chunk", 256)
    ctx.set_option("lanes", lanes)
    try:
        for gamma in (0, 2, 1 << 20):
            ctx.set_option("filter_gamma", gamma)
            for direction in (0, 1):
                ctx.set_option("direction", direction)
                out, ov = ctx.iterativelength(1, ps, pd)
                _cmp_arr("valid", ov, ev)
                _cmp_arr("length", out, eo)
        p = ctx.shortestpath(1, ps, pd, mode=1)
    finally:
        ctx.set_option("tiny_max", 4)
        ctx.set_option("small_max", 64)
        ctx.set_option("med_max", 512)
        ctx.set_option("hub_min", 4096)
        ctx.set_option("hub_chunk", 4096)
        ctx.set_option("lanes", 0)
        ctx.set_option("filter_gamma", 2)
        ctx.set_option("direction", 0)
    assert_paths_equal(p, ep, "paths bins=%s lanes=%d" % (bins, lanes))
    ctx.csr_delete(1)


@pytest.mark.parametrize("direction,beta", [(1, 4), (0, 1), (0, 4), (0, 64)])
@pytest.mark.parametrize("lanes", [64, 512])
def test_direction_switch(ctx, direction, beta, lanes):
    """pull-only, push-only, and mixed schedules (incl. pull->push transitions) give identical answers."""
    scale = 15
    V = 1 << scale
    src, dst = orc.gen_rmat(0x5eed0000 + scale, scale)
    v, e, _ = _build_both(ctx, V, src, dst)
    ps, pd = orc.gen_pairs(12, V, v, 1100)
    ctx.set_option("direction", direction)
    ctx.set_option("push_beta", beta)
    ctx.set_option("push_vbeta", 1 if beta == 1 else 16)
    ctx.set_option("lanes", lanes)
    ctx.set_option("stats", 1)
    ctx.reset_stats()
    try:
        out, ov = ctx.iterativelength(1, ps, pd)
        st = ctx.stats()
    finally:
        ctx.set_option("direction", 0)
        ctx.set_option("push_beta", 32)
        ctx.set_option("push_vbeta", 16)
        ctx.set_option("lanes", 0)
        ctx.set_option("stats", 0)
    eo, ev, est = orc.iterativelength(V, v, e, ps, pd, stats=True)
    _cmp_arr("valid", ov, ev)
    _cmp_arr("length", out, eo)
    if lanes == 512:
        assert st["edges_lane"] == est["edges_lane"] and st["edges_unique"] == est["edges_unique"]
    if direction == 1:
        assert st["push_levels"] == 0
    elif beta == 1:
        assert st["push_levels"] >= st["levels"] - 1
    elif beta == 4:
        assert 0 < st["push_levels"] < st["levels"]
    ctx.csr_delete(1)


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("graph", ["rmat12", "knows"])
def test_shortestpath_bit_identical(ctx, graph, mode):
    if graph == "rmat12":
        V = 1 << 12
        src, dst = orc.gen_rmat(0x5eed0000 + 12, 12)
    else:
        V = 10000
        src, dst = orc.gen_knows(0x1dbc, V, 180000)
    v, e, eids = _build_both(ctx, V, src, dst)
    n = 1300
    ps, pd = orc.gen_pairs(8, V, v, n)
    rng = np.random.default_rng(2)
    sv = (rng.random(n) > 0.02).astype(np.uint8)
    ps[::53] = pd[::53]
    got = ctx.shortestpath(1, ps, pd, sv, None, mode=mode)
    exp = orc.shortestpath(V, v, e, eids, ps, pd, sv, None, mode=mode)
    assert_paths_equal(got, exp, "%s mode %d" % (graph, mode))
    paths = unpack_paths(*got)
    if mode == 0:
        for i in range(0, n, 17):
            if paths[i] is not None and len(paths[i]) > 1:
                check_path_valid(V, v, e, int(ps[i]), int(pd[i]), paths[i])
    ctx.csr_delete(1)


@pytest.mark.parametrize("n", [30, 49, 50, 120])
def test_long_chain_levels(ctx, n):
    """A path graph needs n-1 levels: more than the small-graph kernel keeps level buffers for
    (shortestpath falls back to the general path), and lengths must not care."""
    src = np.arange(n - 1, dtype=np.int64)
    dst = src + 1
    v, e, eids = _build_both(ctx, n, src, dst)
    qs = np.array([0, 0, 1, n - 1, 3], dtype=np.int64)
    qd = np.array([n - 1, n // 2, n - 1, 0, 3], dtype=np.int64)
    out, ov = ctx.iterativelength(1, qs, qd)
    eo, ev = orc.iterativelength(n, v, e, qs, qd)
    assert ov.tolist() == ev.tolist()
    assert lengths_list(out, ov) == lengths_list(eo, ev)
    for mode in (0, 1):
        got = ctx.shortestpath(1, qs, qd, mode=mode)
        exp = orc.shortestpath(n, v, e, eids, qs, qd, mode=mode)
        assert_paths_equal(got, exp, "chain %d mode %d" % (n, mode))
    ctx.csr_delete(1)


def test_config1_ldbc_sf1_shape(ctx):
    """BASELINE config 1: LDBC-SF1-shaped knows graph, CSR build + iterativelength for 2048 pairs."""
    V, E = 10000, 180000
    src, dst = orc.gen_knows(0x1dbc0001, V, E)
    v, e, _ = _build_both(ctx, V, src, dst)
    gv, ge, _ = ctx.csr_download(1)
    _cmp_arr("v", gv, v)
    _cmp_arr("e", ge, e)
    ps, pd = orc.gen_pairs(1, V, v, 2048)
    out, ov = ctx.iterativelength(1, ps, pd)
    _, eo, ev = orc.run_chunks_mt(0, 4, V, v, e, ps, pd)
    _cmp_arr("valid", ov, ev)
    _cmp_arr("length", out, eo)
    ctx.csr_delete(1)


def test_device_resident_variant_and_generators(ctx):
    scale = 13
    V, E = 1 << scale, 16 << scale
    seed = 0x5eed0000 + scale
    d_src = ctx.dev_alloc(E * 8)
    d_dst = ctx.dev_alloc(E * 8)
    ctx.gen_rmat_device(seed, scale, E, d_src, d_dst)
    hs = np.empty(E, dtype=np.int64)
    hd = np.empty(E, dtype=np.int64)
    ctx.d2h(hs, d_src)
    ctx.d2h(hd, d_dst)
    es, ed = orc.gen_rmat(seed, scale)
    _cmp_arr("gen src", hs, es)
    _cmp_arr("gen dst", hd, ed)
    ctx.csr_build_device(2, V, E, d_src, d_dst)
    v, e, _ = orc.csr_build(V, es, ed)
    gv, ge, _ = ctx.csr_download(2)
    _cmp_arr("v", gv, v)
    _cmp_arr("e", ge, e)
    n = 2000
    ps, pd = orc.gen_pairs(4, V, v, n)
    ps[::41] = pd[::41]
    d_ps, d_pd = ctx.dev_alloc(n * 8), ctx.dev_alloc(n * 8)
    d_out, d_ov = ctx.dev_alloc(n * 8), ctx.dev_alloc(n)
    ctx.h2d(d_ps, ps)
    ctx.h2d(d_pd, pd)
    ctx.iterativelength_device(2, n, d_ps, d_pd, d_out, d_ov)
    out = np.empty(n, dtype=np.int64)
    ov = np.empty(n, dtype=np.uint8)
    ctx.d2h(out, d_out)
    ctx.d2h(ov, d_ov)
    eo, ev = orc.iterativelength(V, v, e, ps, pd)
    _cmp_arr("valid", ov, ev)
    _cmp_arr("length", out, eo)
    # knows generator parity
    KV, KE = 7000, 50000
    mult, add = orc.knows_params(KV)
    ctx.gen_knows_device(9, KV, KE, mult, add, d_src, d_dst)
    ks = np.empty(KE, dtype=np.int64)
    kd = np.empty(KE, dtype=np.int64)
    ctx.d2h(ks, d_src)
    ctx.d2h(kd, d_dst)
    oks, okd = orc.gen_knows(9, KV, KE)
    _cmp_arr("knows src", ks, oks)
    _cmp_arr("knows dst", kd, okd)
    for p in (d_src, d_dst, d_ps, d_pd, d_out, d_ov):
        ctx.dev_free(p)
    ctx.csr_delete(2)


def test_large_scale_properties(ctx):
    """R-MAT scale 20 (1M vertices, 16M edges): sizes where the oracle is slow.  Checks
    size-independent properties: CSR checksums vs a numpy stable sort, lane-width independence,
    path validity and |path| == iterativelength + 1, sampled textbook BFS."""
    scale = 20
    V, E = 1 << scale, 16 << scale
    src, dst = orc.gen_rmat(0x5eed0000 + scale, scale)
    ctx.csr_build(9, V, src, dst)
    gv, ge, _ = ctx.csr_download(9)
    order = np.argsort(src, kind="stable")
    _cmp_arr("e (stable sort)", ge, dst[order])
    cnt = np.bincount(src, minlength=V)
    _cmp_arr("v", gv, np.concatenate([[0], np.cumsum(cnt), [E]]))
    n = 1024
    ps, pd = orc.gen_pairs(2, V, gv, n)
    out512, ov512 = ctx.iterativelength(9, ps, pd)
    ctx.set_option("lanes", 64)
    try:
        out64, ov64 = ctx.iterativelength(9, ps[:256], pd[:256])
    finally:
        ctx.set_option("lanes", 0)
    _cmp_arr("lane-width independence", out64, out512[:256])
    for i in range(0, n, 128):
        dist = orc.bfs_simple(V, gv, ge, int(ps[i]))
        assert (int(out512[i]) if ov512[i] else -1) == int(dist[pd[i]])
    paths = unpack_paths(*ctx.shortestpath(9, ps[:512], pd[:512]))
    for i in range(512):
        if ov512[i]:
            assert paths[i] is not None and len(paths[i]) == out512[i] + 1
            if i % 16 == 0:
                check_path_valid(V, gv, ge, int(ps[i]), int(pd[i]), paths[i])
        else:
            assert paths[i] is None
    ctx.csr_delete(9)


@pytest.mark.gpu
def test_sharded_query_gpu(ctx):
    """The multi-GPU host wrapper around the real per-GPU query (single rank = whole set)."""
    import pgq_b200
    scale = 12
    V = 1 << scale
    src, dst = orc.gen_rmat(0x5eed0000 + scale, scale)
    v, e, _ = _build_both(ctx, V, src, dst)
    ps, pd = orc.gen_pairs(8, V, v, 1500)
    outs, valids = [], []
    for r in range(3):  # what three ranks would each compute, stitched by hand
        lo, hi = pgq_b200.shard_rows(len(ps), 3, r)
        o, ok = pgq_b200.sharded_query(lambda s, d: ctx.iterativelength(1, s, d), None, hi - lo, 1, 0, ps[lo:hi],
                                       pd[lo:hi])
        outs.append(o)
        valids.append(ok)
    eo, ev = orc.iterativelength(V, v, e, ps, pd)
    _cmp_arr("valid", np.concatenate(valids), ev)
    _cmp_arr("length", np.concatenate(outs), eo)
    ctx.csr_delete(1)


@pytest.mark.gpu
@pytest.mark.parametrize("lanes", [64, 512])
def test_partitioned_path_world1(ctx, lanes):
    """1-D vertex-partitioned MS-BFS (BASELINE configs[4]) degenerated to one rank: the slice build
    (owner filter, renumbering, stable sort), the replicated-frontier level loop and the
    next[dst]-based destination check give the oracle's lengths / reachability.  The multi-rank run
    of the same code is tools/part_check.py --check under torchrun (2 GPUs)."""
    scale = 13
    V = 1 << scale
    src, dst = orc.gen_rmat(0x5eed0000 + scale, scale)
    E = len(src)
    v, e, _ = orc.csr_build(V, src, dst)
    ps, pd = orc.gen_pairs(21, V, v, 1300)
    sv = np.ones(len(ps), dtype=np.uint8)
    sv[::53] = 0
    pd[7::61] = ps[7::61]
    d_src, d_dst = ctx.dev_alloc(E * 8), ctx.dev_alloc(E * 8)
    ctx.h2d(d_src, src)
    ctx.h2d(d_dst, dst)
    ctx.part_init(0, 1)
    ctx.part_csr_create(7, V)
    half = E // 2
    ctx.part_csr_add_edges_device(7, d_src, d_dst, half)  # two chunks, arrival order
    ctx.part_csr_add_edges_device(7, d_src + half * 8, d_dst + half * 8, E - half)
    ctx.part_csr_finalize(7)
    ctx.set_option("lanes", lanes)
    try:
        out, ov = ctx.part_iterativelength(7, ps, pd, src_valid=sv)
        r, rv = ctx.part_reachability(7, ps, pd, src_valid=sv)
        with pytest.raises(Exception):
            ctx.shortestpath(7, ps[:4], pd[:4])  # not offered on the partitioned path
    finally:
        ctx.set_option("lanes", 0)
    eo, ev = orc.iterativelength(V, v, e, ps, pd, src_valid=sv)
    _cmp_arr("valid", ov, ev)
    _cmp_arr("length", out, eo)
    er, erv = orc.reachability(V, v, e, ps, pd, src_valid=sv)
    _cmp_arr("reach valid", rv, erv)
    _cmp_arr("reach", r, er)
    ctx.dev_free(d_src)
    ctx.dev_free(d_dst)
    ctx.csr_delete(7)


# ---- weighted next row (rule R10) -----------------------------------------------

def _weighted_case(ctx, V, src, dst, w, n_pairs, seed):
    v, e, eids = _build_both(ctx, V, src, dst)
    wc = orc.csr_weights(V, v, src, w)
    ctx.csr_set_weights(1, w, chunk=1024)  # chunks as the engine would hand them over
    _cmp_arr("weights in CSR order", ctx.csr_download_weights(1), wc)
    ps, pd = orc.gen_pairs(seed, V, v, n_pairs)
    rng = np.random.default_rng(seed)
    sv = (rng.random(n_pairs) > 0.03).astype(np.uint8)
    dv = (rng.random(n_pairs) > 0.03).astype(np.uint8)
    ps[::41] = pd[::41]
    got, gv = ctx.cheapest_path_length(1, ps, pd, sv, dv)
    exp, ev, _ = orc.cheapest_path_length(V, v, e, wc, ps, pd, sv, dv)
    _cmp_arr("cheapest_path_length validity", gv, ev)
    assert got.dtype == exp.dtype
    assert got.tobytes() == exp.tobytes(), "cheapest_path_length values differ (bit-exact expected)"
    return v, e, wc, ps, pd, got, gv


@pytest.mark.parametrize("graph", ["rmat12", "knows"])
def test_cheapest_path_length_int64(ctx, graph):
    if graph == "rmat12":
        V = 1 << 12
        src, dst = orc.gen_rmat(0x5eed0000 + 12, 12)
    else:
        V = 3000
        src, dst = orc.gen_knows(0x1dbc, V, 50000)
    w = np.random.default_rng(11).integers(1, 1000, size=len(src)).astype(np.int64)
    v, e, wc, ps, pd, got, gv = _weighted_case(ctx, V, src, dst, w, 200, 5)
    # same answers when almost every vertex goes through the block-per-vertex kernel
    ctx.set_option("small_max", 2)
    ctx.set_option("med_max", 1)  # clamped to small_max internally
    try:
        again, av = ctx.cheapest_path_length(1, ps, pd)
        ref, rv, _ = orc.cheapest_path_length(V, v, e, wc, ps, pd)
        assert av.tolist() == rv.tolist() and again.tobytes() == ref.tobytes()
    finally:
        ctx.set_option("small_max", 64)
        ctx.set_option("med_max", 512)
    for i in range(0, 200, 23):  # independent check: binary-heap Dijkstra
        if ps[i] == pd[i]:
            continue
        d = orc.dijkstra(V, v, e, wc, int(ps[i]))[pd[i]]
        if gv[i]:
            assert got[i] == d
    ctx.csr_delete(1)


def test_cheapest_path_length_double_and_negative(ctx):
    V = 1 << 11
    src, dst = orc.gen_rmat(0x5eed0000 + 11, 11)
    w = np.random.default_rng(12).random(len(src)) * 10.0 + 0.001
    _weighted_case(ctx, V, src, dst, w, 150, 6)
    ctx.csr_delete(1)
    # negative weights on a DAG (edges only from lower to higher ids): no cycle, Bellman-Ford is exact
    lo, hi = np.minimum(src, dst), np.maximum(src, dst)
    keep = lo != hi
    lo, hi = lo[keep], hi[keep]
    w = np.random.default_rng(13).integers(-50, 100, size=len(lo)).astype(np.int64)
    _weighted_case(ctx, V, lo, hi, w, 150, 7)
    ctx.csr_delete(1)


@pytest.mark.parametrize("wcase", ["fits16", "overflow16", "huge", "zeros"])
def test_cheapest_path_length_narrow_rows_and_bound(ctx, wcase):
    """Non-negative int64 weights run on 16-bit distance rows with the destination's distance as a bound; a batch
    whose distances do not fit is redone on 32-bit rows, and on 8-byte rows if they do not fit those either.  Same
    bytes as the oracle in every case, and the same as with the shortcuts switched off."""
    V = 2500
    src, dst = orc.gen_knows(0x1dbc + 7, V, 30000)
    rng = np.random.default_rng(21)
    if wcase == "fits16":
        w = rng.integers(1, 101, size=len(src))
    elif wcase == "overflow16":
        w = rng.integers(20000, 60000, size=len(src))  # two hops already pass 65535
    elif wcase == "huge":
        w = np.where(rng.random(len(src)) < 0.3, rng.integers(1 << 31, 1 << 40, size=len(src)),
                     rng.integers(0, 50, size=len(src)))
    else:
        w = np.where(rng.random(len(src)) < 0.5, 0, rng.integers(0, 3, size=len(src)))
    w = w.astype(np.int64)
    v, e, wc, ps, pd, got, gv = _weighted_case(ctx, V, src, dst, w, 300, 9)
    ctx.reset_stats()
    ctx.cheapest_path_length(1, ps, pd)
    st = ctx.stats()
    if wcase in ("fits16", "zeros"):
        assert st["sssp_narrow_batches"] == st["batches"] > 0 and st["sssp_redo_batches"] == 0, st
    if wcase == "overflow16":  # too long for 14 bits, fine for 30: redone once, on 32-bit rows
        assert st["sssp_redo_batches"] == st["sssp_narrow32_batches"] > 0, st
        assert st["sssp_narrow_batches"] + st["sssp_narrow32_batches"] == st["batches"] - st["sssp_redo_batches"], st
    if wcase == "huge":  # weights past 2^30: some batches end on 8-byte rows after two redos
        assert st["sssp_redo_batches"] > st["sssp_narrow32_batches"], st
    try:
        for narrow, prune, slots in ((0, 1, 4), (1, 0, 1), (0, 0, 2), (2, 1, 3)):
            ctx.set_option("sssp_narrow", narrow)
            ctx.set_option("sssp_prune", prune)
            ctx.set_option("sssp_slots", slots)
            again, av = ctx.cheapest_path_length(1, ps, pd)
            ref, rv, _ = orc.cheapest_path_length(V, v, e, wc, ps, pd)
            assert av.tolist() == rv.tolist() and again.tobytes() == ref.tobytes(), (narrow, prune, slots)
    finally:
        ctx.set_option("sssp_narrow", 1)
        ctx.set_option("sssp_prune", 1)
        ctx.set_option("sssp_slots", 4)
    ctx.csr_delete(1)


@pytest.mark.parametrize("kind", ["narrow", "narrow32", "wide", "double", "negative"])
def test_cheapest_path_length_hub_chunks(ctx, kind):
    """In-lists above hub_min are relaxed chunk by chunk, the chunks' minima meeting through atomic minima (uint32,
    int64, and compare-and-swap for double): force that path on a small graph."""
    V = 1 << 11
    src, dst = orc.gen_rmat(0x5eed0000 + 11, 11)
    rng = np.random.default_rng(31)
    if kind == "double":
        w = rng.random(len(src)) * 7.0
    elif kind == "negative":
        lo, hi = np.minimum(src, dst), np.maximum(src, dst)
        keep = lo != hi
        src, dst = lo[keep], hi[keep]
        w = rng.integers(-20, 60, size=len(src)).astype(np.int64)
    else:
        w = rng.integers(0, 200, size=len(src)).astype(np.int64)
    ctx.set_option("hub_min", 8)
    ctx.set_option("hub_chunk", 256)
    ctx.set_option("sssp_narrow", {"wide": 0, "narrow32": 2}.get(kind, 1))
    try:
        _weighted_case(ctx, V, src, dst, w, 260, 17)
        ctx.csr_delete(1)
    finally:
        ctx.set_option("hub_min", 4096)
        ctx.set_option("hub_chunk", 4096)
        ctx.set_option("sssp_narrow", 1)


def test_cheapest_path_length_whole_numbered_doubles(ctx):
    """DOUBLE weights that are all non-negative whole numbers take the integer rows (sums of such doubles are exact, so
    the bits are the oracle's); one fractional weight, and the same graph runs on 8-byte double rows."""
    V = 2500
    src, dst = orc.gen_knows(0x1dbc + 9, V, 30000)
    w = np.random.default_rng(41).integers(0, 300, size=len(src)).astype(np.float64)
    _weighted_case(ctx, V, src, dst, w, 300, 11)
    ctx.reset_stats()
    v, e, _ = orc.csr_build(V, src, dst, want_eids=False)
    ps, pd = orc.gen_pairs(3, V, v, 128)
    ctx.cheapest_path_length(1, ps, pd)
    st = ctx.stats()
    assert st["sssp_narrow_batches"] == st["batches"] == 2, st
    ctx.csr_delete(1)
    w[17] = 2.5
    _weighted_case(ctx, V, src, dst, w, 300, 11)
    ctx.reset_stats()
    ctx.cheapest_path_length(1, ps, pd)
    st = ctx.stats()
    assert st["sssp_narrow_batches"] == 0 and st["sssp_narrow32_batches"] == 0, st
    ctx.csr_delete(1)
    w[17] = 3.0e6  # whole, but past the bound that keeps every sum exact: double rows again
    _weighted_case(ctx, V, src, dst, w, 300, 11)
    ctx.reset_stats()
    ctx.cheapest_path_length(1, ps, pd)
    assert ctx.stats()["sssp_narrow_batches"] == 0
    ctx.csr_delete(1)


def test_cheapest_path_length_errors(ctx):
    src = np.array([0, 1, 2], dtype=np.int64)
    dst = np.array([1, 2, 0], dtype=np.int64)
    ctx.csr_build(1, 3, src, dst, rowids=True)
    import pgq_b200
    with pytest.raises(pgq_b200.PgqError):  # no weights yet
        ctx.cheapest_path_length(1, [0], [2])
    ctx.csr_set_weights(1, np.array([1, -5, 1], dtype=np.int64))
    with pytest.raises(pgq_b200.PgqError, match="negative cycle"):
        ctx.cheapest_path_length(1, [0], [2])
    ctx.csr_set_weights(1, np.array([1.5, 2.25, 1.0]))
    out, ov = ctx.cheapest_path_length(1, [0, 2, 1], [2, 1, 1])
    assert ov.tolist() == [1, 1, 1] and out.tolist() == [3.75, 2.5, 0.0]
    ctx.csr_delete(1)
    ctx.csr_build(1, 3, src, dst)  # no row tracking
    with pytest.raises(pgq_b200.PgqError, match="row tracking"):
        ctx.csr_set_weights(1, np.array([1, 1, 1], dtype=np.int64))
    ctx.csr_delete(1)


@pytest.mark.parametrize("streams", [1, 3, 8])
def test_batches_in_flight(ctx, streams):
    """Batches of one call run concurrently on helper streams (option `streams`); lengths,
    reachability and paths (stitched in row order, R9) must not depend on it."""
    scale = 13
    V = 1 << scale
    src, dst = orc.gen_rmat(0x5eed0000 + scale, scale)
    v, e, eids = _build_both(ctx, V, src, dst)
    n = 1024
    ps, pd = orc.gen_pairs(11, V, v, n)
    sv = (np.random.default_rng(3).random(n) > 0.02).astype(np.uint8)
    ps[::41] = pd[::41]
    ctx.set_option("lanes", 64)  # 16 batches per call
    ctx.set_option("streams", streams)
    try:
        out, ov = ctx.iterativelength(1, ps, pd, sv, None)
        got = ctx.shortestpath(1, ps, pd, sv, None, mode=1)
        st = ctx.stats()
    finally:
        ctx.set_option("lanes", 0)
        ctx.set_option("streams", 4)
    eo, ev = orc.iterativelength(V, v, e, ps, pd, sv, None)
    _cmp_arr("valid", ov, ev)
    _cmp_arr("length", out, eo)
    exp = orc.shortestpath(V, v, e, eids, ps, pd, sv, None, mode=1)
    assert_paths_equal(got, exp, "streams %d" % streams)
    assert st["batches"] >= 16
    ctx.csr_delete(1)


# ---- round 2: device-side level loop, done bitmap, level-history budget, multi-GPU context ----------------
def _rmat_case(ctx, scale, n, seed=21):
    V = 1 << scale
    src, dst = orc.gen_rmat(0x5eed0000 + scale, scale)
    v, e, eids = _build_both(ctx, V, src, dst)
    ps, pd = orc.gen_pairs(seed, V, v, n)
    return V, v, e, eids, ps, pd


def test_device_loop_has_no_host_sync_and_equals_host_loop(ctx):
    """The level loop runs as one CUDA graph (WHILE + SWITCH nodes) per batch: between the init kernel and
    the result copy the host never waits (stats.host_syncs == 0), and the answers equal the host-driven
    loop's and the oracle's."""
    V, v, e, eids, ps, pd = _rmat_case(ctx, 14, 1500)
    eo, ev = orc.iterativelength(V, v, e, ps, pd)
    res = {}
    try:
        ctx.set_option("small_path", 0)
        for loop in (1, 0):
            ctx.set_option("loop", loop)
            ctx.reset_stats()
            out, ov = ctx.iterativelength(1, ps, pd)
            st = ctx.stats()
            _cmp_arr("valid loop=%d" % loop, ov, ev)
            _cmp_arr("length loop=%d" % loop, out, eo)
            res[loop] = st
    finally:
        ctx.set_option("loop", 1)
    assert res[1]["host_syncs"] == 0, "device loop fell back to the host loop: %s" % res[1]
    assert res[0]["host_syncs"] >= res[0]["levels"] > 0
    assert res[1]["levels"] == res[0]["levels"] and res[1]["push_levels"] == res[0]["push_levels"]
    assert res[1]["batches"] == res[0]["batches"] == 3
    ctx.csr_delete(1)


def test_loop_graph_is_built_only_where_it_pays():
    """Default thresholds (own context): a small call on a fresh CSR takes the host-driven loop (building the loop
    graph would cost more than the call), the fifth call on the same CSR and any call of >= 8 batches take the
    device loop.  Same answers either way."""
    import pgq_b200
    c = pgq_b200.Context(0)
    try:
        c.set_option("small_path", 0)
        scale = 13
        V = 1 << scale
        src, dst = orc.gen_rmat(0x5eed0000 + scale, scale)
        c.csr_build(1, V, src, dst)
        v, e, _ = orc.csr_build(V, src, dst, want_eids=False)
        ps, pd = orc.gen_pairs(2, V, v, 600)
        eo, ev = orc.iterativelength(V, v, e, ps, pd)
        syncs = []
        for _ in range(6):
            c.reset_stats()
            out, ov = c.iterativelength(1, ps, pd)
            _cmp_arr("valid", ov, ev)
            _cmp_arr("length", out, eo)
            syncs.append(c.stats()["host_syncs"])
        assert all(x > 0 for x in syncs[:4]) and syncs[4] == 0 and syncs[5] == 0, syncs
        c.csr_build(2, V, src, dst)  # fresh CSR, 10 batches of 64 lanes in one call
        c.set_option("lanes", 64)
        c.reset_stats()
        out, ov = c.iterativelength(2, ps, pd)
        _cmp_arr("valid 64", ov, ev)
        _cmp_arr("length 64", out, eo)
        assert c.stats()["host_syncs"] == 0 and c.stats()["batches"] == 10
    finally:
        c.close()


@pytest.mark.parametrize("done,l2p", [(0, 0), (1, 1)])
def test_done_bitmap_and_l2_window_do_not_change_results(ctx, done, l2p):
    V, v, e, eids, ps, pd = _rmat_case(ctx, 13, 700, seed=5)
    ctx.set_option("done_bitmap", done)
    ctx.set_option("l2_persist", l2p)
    try:
        out, ov = ctx.iterativelength(1, ps, pd)
        got = ctx.shortestpath(1, ps[:300], pd[:300], mode=1)
    finally:
        ctx.set_option("done_bitmap", 1)
        ctx.set_option("l2_persist", 0)
    eo, ev = orc.iterativelength(V, v, e, ps, pd)
    _cmp_arr("valid", ov, ev)
    _cmp_arr("length", out, eo)
    assert_paths_equal(got, orc.shortestpath(V, v, e, eids, ps[:300], pd[:300], mode=1), "done=%d" % done)
    ctx.csr_delete(1)


def test_level_history_budget(ctx):
    """shortestpath keeps one frontier buffer per level.  When a batch's history would exceed its budget the
    batch is redone with 64-lane words (8x fewer bytes per level), beyond that the call fails with a message
    naming the option -- never a raw allocation error; results are unchanged by the narrowing."""
    import pgq_b200
    n = 8000
    src = np.arange(n - 1, dtype=np.int64)
    dst = src + 1
    v, e, eids = _build_both(ctx, n, src, dst)
    rng = np.random.default_rng(9)
    qs = rng.integers(0, 200, 300).astype(np.int64)
    qd = (qs + rng.integers(1, 120, 300)).astype(np.int64)  # up to 120 levels
    qs[7] = qd[7]
    exp = orc.shortestpath(n, v, e, eids, qs, qd, mode=1)
    ctx.set_option("lanes", 512)
    try:
        ctx.set_option("hist_budget_mb", 16)  # 512 lanes: 512 KB per level -> 32 levels; 64 lanes: 256 levels
        got = ctx.shortestpath(1, qs, qd, mode=1)
        assert_paths_equal(got, exp, "narrowed history")
        ctx.set_option("hist_budget_mb", 1)  # 64 lanes: 16 levels -- not enough for 120
        with pytest.raises(pgq_b200.PgqError, match="hist_budget_mb"):
            ctx.shortestpath(1, qs, qd, mode=1)
        ctx.set_option("hist_budget_mb", 0)
        got = ctx.shortestpath(1, qs, qd, mode=0)
        assert_paths_equal(got, orc.shortestpath(n, v, e, eids, qs, qd, mode=0), "default budget")
    finally:
        ctx.set_option("hist_budget_mb", 0)
        ctx.set_option("lanes", 0)
    ctx.csr_delete(1)


def test_model_byte_counters(ctx):
    """Instrumented pass (stats=1): what the level kernels touched, the numerator of bench.py's roofline."""
    V, v, e, eids, ps, pd = _rmat_case(ctx, 14, 512, seed=2)
    ctx.set_option("small_path", 0)
    ctx.set_option("stats", 1)
    try:
        ctx.reset_stats()
        out, ov = ctx.iterativelength(1, ps, pd)
        st = ctx.stats()
    finally:
        ctx.set_option("stats", 0)
    eo, ev, est = orc.iterativelength(V, v, e, ps, pd, stats=True)
    _cmp_arr("length", out, eo)
    assert st["edges_lane"] == est["edges_lane"]
    assert st["k_listed"] > 0 and st["k_open"] > 0 and st["k_new"] > 0
    assert st["k_scanned"] >= st["k_gathered"] > 0
    assert st["k_open"] <= st["k_listed"]
    E = len(e)
    assert st["k_scanned"] <= E * st["levels"]
    lower = st["k_gathered"] * 64 + st["k_scanned"] * 4
    assert lower < st["model_bytes"] < st["algo_bytes"] * 4
    ctx.csr_delete(1)


def test_multi_gpu_context_one_process(ctx):
    """pgq_init_devices: ONE process, the batches of one call dealt over several GPUs (CSR replicated over
    NVLink on first use): same answers as one device, for lengths, reachability and paths."""
    import pgq_b200
    from conftest import gpu_count
    ng = gpu_count()
    if ng < 2:
        pytest.skip("needs >= 2 GPUs")
    V, v, e, eids, ps, pd = _rmat_case(ctx, 14, 4096, seed=8)
    ctx.csr_delete(1)
    src, dst = orc.gen_rmat(0x5eed0000 + 14, 14)
    m = pgq_b200.Context(devices=list(range(min(ng, 4))))
    try:
        m.set_option("small_path", 0)
        m.set_option("lanes", 64)  # 64 batches
        m.csr_build(3, V, src, dst, rowids=True)
        out, ov = m.iterativelength(3, ps, pd)
        r, rv = m.reachability(3, ps, pd)
        got = m.shortestpath(3, ps[:1000], pd[:1000], mode=1)
        st = m.stats()
    finally:
        m.close()
    eo, ev = orc.iterativelength(V, v, e, ps, pd)
    _cmp_arr("valid", ov, ev)
    _cmp_arr("length", out, eo)
    assert r.tolist() == [int(x) for x in ev]
    assert_paths_equal(got, orc.shortestpath(V, v, e, eids, ps[:1000], pd[:1000], mode=1), "multi-gpu paths")
    assert st["batches"] >= 64
