"""CPU suite: pins the oracle against the hand-checked golden graphs, a textbook BFS and the
fixtures produced by the reference DuckDB's recursive-CTE BFS (tests/golden/)."""
import json
import os

import numpy as np
import pytest

import orc
from golden_cases import CASES, CSR_EXPECT
from helpers import lengths_list, unpack_paths, check_path_valid

GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("name", sorted(CASES))
def test_golden_case(name):
    c = CASES[name]
    src = [a for a, _ in c["edges"]]
    dst = [b for _, b in c["edges"]]
    v, e, eid = orc.csr_build(c["V"], src, dst)
    if name in CSR_EXPECT:
        assert v.tolist() == CSR_EXPECT[name]["v"]
        assert e.tolist() == CSR_EXPECT[name]["e"]
        assert eid.tolist() == CSR_EXPECT[name]["eid"]
    qs = [a for a, _ in c["q"]]
    qd = [b for _, b in c["q"]]
    out, ov = orc.iterativelength(c["V"], v, e, qs, qd)
    assert lengths_list(out, ov) == c["lengths"]
    r, rv = orc.reachability(c["V"], v, e, qs, qd)
    assert rv.tolist() == [1] * len(qs) and r.tolist() == c["reach"]
    for faithful in (False, True):
        p = orc.shortestpath(c["V"], v, e, eid, qs, qd, mode=0, faithful=faithful)
        assert unpack_paths(*p) == c["paths"]
        pe = orc.shortestpath(c["V"], v, e, eid, qs, qd, mode=1, faithful=faithful)
        assert unpack_paths(*pe) == c["paths_e"]


def test_null_rows_G8():
    c = CASES["G1_path"]
    src = [a for a, _ in c["edges"]]
    dst = [b for _, b in c["edges"]]
    v, e, eid = orc.csr_build(4, src, dst)
    qs, qd = [0, 0, 1, 2], [3, 3, 3, 2]
    sv, dv = [1, 0, 1, 1], [1, 1, 0, 1]
    out, ov = orc.iterativelength(4, v, e, qs, qd, sv, dv)
    assert lengths_list(out, ov) == [3, None, None, 0]
    r, rv = orc.reachability(4, v, e, qs, qd, sv, dv)
    assert rv.tolist() == [1, 0, 0, 1] and r.tolist()[0] == 1
    p = unpack_paths(*orc.shortestpath(4, v, e, eid, qs, qd, sv, dv))
    assert p == [[0, 1, 2, 3], None, None, [2]]


def test_cte_golden_upper_bound():
    """Upper-bounded hop distances produced by the reference DuckDB's recursive CTE (tests/golden/
    make_cte_golden2.py) on a knows-shaped graph: pins iterativelength, reachability and the upper bound of R11."""
    g = json.load(open(os.path.join(GOLD, "cte_knows_bounded.json")))
    v, e, _ = orc.csr_build(g["V"], g["src"], g["dst"])
    top = str(max(g["uppers"]))
    out, ov = orc.iterativelength(g["V"], v, e, g["q_src"], g["q_dst"])
    assert lengths_list(out, ov) == g["expected"][top]
    r, rv = orc.reachability(g["V"], v, e, g["q_src"], g["q_dst"])
    assert r.tolist() == [int(x is not None) for x in g["expected"][top]]
    for u in g["uppers"]:
        out, ov = orc.iterativelength_bounded(g["V"], v, e, g["q_src"], g["q_dst"], 0, u)
        assert lengths_list(out, ov) == g["expected"][str(u)], u


def test_bounded_lengths_R11():
    """6-argument iterativelength: hand-checked on the path graph, then the defining property on R-MAT:
    bounded(l, u) = full where l <= full <= u, NULL elsewhere."""
    c = CASES["G1_path"]  # 0 -> 1 -> 2 -> 3
    v, e, _ = orc.csr_build(4, [a for a, _ in c["edges"]], [b for _, b in c["edges"]])
    qs, qd = [0, 0, 0, 1, 2, 3], [3, 2, 1, 3, 2, 0]
    assert lengths_list(*orc.iterativelength_bounded(4, v, e, qs, qd, 0, 2)) == [None, 2, 1, 2, 0, None]
    assert lengths_list(*orc.iterativelength_bounded(4, v, e, qs, qd, 0, 3)) == [3, 2, 1, 2, 0, None]
    assert lengths_list(*orc.iterativelength_bounded(4, v, e, qs, qd, 2, 3)) == [3, 2, None, 2, None, None]
    assert lengths_list(*orc.iterativelength_bounded(4, v, e, qs, qd, 1, 1)) == [None, None, 1, None, None, None]
    assert lengths_list(*orc.iterativelength_bounded(4, v, e, qs, qd, 0, 0)) == [None, None, None, None, 0, None]
    assert lengths_list(*orc.iterativelength_bounded(4, v, e, qs, qd, 0, -1)) == [3, 2, 1, 2, 0, None]
    s, d = orc.gen_rmat(21, 11)
    V = 1 << 11
    v, e, _ = orc.csr_build(V, s, d)
    ps, pd = orc.gen_pairs(9, V, v, 1300)
    ps[::53] = pd[::53]
    sv = (np.arange(1300) % 41 != 0).astype(np.uint8)
    full, fv = orc.iterativelength(V, v, e, ps, pd, sv)
    for lo, hi in [(0, 1), (0, 3), (2, 4), (3, 3), (1, 100), (5, 2)]:
        out, ov = orc.iterativelength_bounded(V, v, e, ps, pd, lo, hi, sv)
        ok = (fv == 1) & (full >= lo) & (full <= hi)
        assert np.array_equal(ov, ok.astype(np.uint8)), (lo, hi)
        assert np.array_equal(out, np.where(ok, full, -1)), (lo, hi)


def test_batches_G9_G10():
    """> 512 and > 1024 rows (lane reuse), duplicate pairs inside one batch."""
    s, d = orc.gen_rmat(11, 10)
    V = 1 << 10
    v, e, eid = orc.csr_build(V, s, d)
    ps, pd = orc.gen_pairs(3, V, v, 1500)
    ps[100:140] = ps[100]  # duplicates
    pd[100:140] = pd[100]
    ps[700] = pd[700]      # trivial row in the middle
    out, ov = orc.iterativelength(V, v, e, ps, pd)
    assert len(set(out[100:140].tolist())) == 1
    for i in range(0, 1500, 13):
        dist = orc.bfs_simple(V, v, e, int(ps[i]))
        assert (int(out[i]) if ov[i] else -1) == int(dist[pd[i]])
    # chunked (<=1024 rows per call, like DuckDB) == one big call
    _, out2, ov2 = orc.run_chunks_mt(0, 3, V, v, e, ps, pd)
    assert np.array_equal(out, out2) and np.array_equal(ov, ov2)
    off, ln, pv, vals = orc.shortestpath(V, v, e, eid, ps, pd)
    assert np.array_equal(pv, ov)
    paths = unpack_paths(off, ln, pv, vals)
    for i in range(0, 1500, 7):
        if paths[i] is not None:
            assert len(paths[i]) == out[i] + 1
            check_path_valid(V, v, e, int(ps[i]), int(pd[i]), paths[i])
    # offsets are a running total in row order (R9)
    assert np.array_equal(off, np.concatenate([[0], np.cumsum(ln)[:-1]]).astype(np.uint64))


def test_faithful_parent_loop_equals_fast():
    s, d = orc.gen_knows(5, 300, 3000)
    v, e, eid = orc.csr_build(300, s, d)
    ps, pd = orc.gen_pairs(9, 300, v, 600)
    a = orc.shortestpath(300, v, e, eid, ps, pd, mode=1, faithful=True)
    b = orc.shortestpath(300, v, e, eid, ps, pd, mode=1, faithful=False)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


def test_generators_deterministic_and_shaped():
    s1, d1 = orc.gen_rmat(0x5eed0000 + 12, 12)
    s2, d2 = orc.gen_rmat(0x5eed0000 + 12, 12, nthreads=3)
    assert np.array_equal(s1, s2) and np.array_equal(d1, d2)
    a, b = orc.gen_rmat(0x5eed0000 + 12, 12, row0=1000, nrows=50)
    assert np.array_equal(a, s1[1000:1050]) and np.array_equal(b, d1[1000:1050])
    assert s1.max() < (1 << 12) and s1.min() >= 0 and len(s1) == 16 << 12
    # quadrant a dominates: most edges land in the low-id corner
    assert (s1 < (1 << 11)).mean() > 0.7
    ks, kd = orc.gen_knows(7, 10000, 180000)
    assert np.array_equal(ks[:90000], kd[90000:]) and np.array_equal(kd[:90000], ks[90000:])
    assert (ks != kd).all()


def test_cte_golden_lengths():
    """Hop distances computed by the REFERENCE DuckDB (recursive CTE of
    benchmark/ldbc/queries/bi-10-shortestpath.sql:8-31) == oracle iterativelength."""
    path = os.path.join(GOLD, "cte_lengths.json")
    if not os.path.exists(path):
        pytest.skip("fixture missing")
    g = json.load(open(path))
    v, e, _ = orc.csr_build(g["V"], g["src"], g["dst"])
    out, ov = orc.iterativelength(g["V"], v, e, g["q_src"], g["q_dst"])
    assert lengths_list(out, ov) == g["expected"]
