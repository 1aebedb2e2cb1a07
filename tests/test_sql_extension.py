"""The drop-in boundary through SQL: the DuckDB extension (duckdb-pgq.old_b200/ext) LOADed into the
reference DuckDB (libduckdb.so built from /root/reference, baseline/_ref/duckdb), driven by SQL text
the way DuckPGQ's rewriter calls the functions (SURVEY.md §8(f) f3 template), checked against the
oracle.  Mirrors /root/reference/test/extension/load_extension.test (LOAD + call)."""
import os
import subprocess

import numpy as np
import pytest

import orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXT = os.path.join(ROOT, "duckdb-pgq.old_b200", "ext", "pgq_b200.duckdb_extension")
DRIVER = os.path.join(ROOT, "duckdb-pgq.old_b200", "ext", "pgq_sql")
HOSTLIB = os.path.join(ROOT, "baseline", "_ref", "duckdb", "libduckdb.so")

needs_host = pytest.mark.skipif(not (os.path.exists(EXT) and os.path.exists(DRIVER) and os.path.exists(HOSTLIB)),
                                reason="reference DuckDB host / extension not built (build container only)")


def run_sql(sql):
    """-> list of (kind, payload): ('result', [rows as lists of str]) or ('error', message)."""
    out = subprocess.run([DRIVER], input=sql, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    res, cur = [], None
    for line in out.stdout.splitlines():
        if line.startswith("-- result "):
            cur = []
            res.append(("result", cur))
        elif line.startswith("-- error "):
            res.append(("error", line.split(":", 1)[1].strip()))
            cur = None
        elif line.startswith("-- done "):
            cur = None
        elif cur is not None:
            cur.append(line)
    return res


@needs_host
def test_extension_loads_and_fails_loudly_without_gpu():
    """LOAD registers every function; with no GPU a call raises (no CPU fallback), never crashes."""
    import torch
    sql = "LOAD '%s';\n" % EXT
    sql += "SELECT name FROM pragma_functions() WHERE name IN ('create_csr_vertex'," \
           "'create_csr_edge','iterativelength','reachability','shortestpath','shortestpath_with_edges'," \
           "'delete_csr','cheapest_path_length','cheapest_path_length_double') GROUP BY name ORDER BY name;\n"
    sql += "SELECT iterativelength(0, CAST(10 AS BIGINT), CAST(1 AS BIGINT), CAST(2 AS BIGINT));\n"
    # every overload binds (the error, if any, comes from the call, not from the binder)
    sql += "SELECT iterativelength(0, CAST(10 AS BIGINT), CAST(1 AS BIGINT), CAST(2 AS BIGINT), CAST(0 AS BIGINT), CAST(3 AS BIGINT));\n"
    sql += "SELECT reachability(0, TRUE, CAST(10 AS BIGINT), CAST(1 AS BIGINT), CAST(2 AS BIGINT));\n"
    sql += "SELECT create_csr_edge(0, CAST(10 AS BIGINT), CAST(1 AS BIGINT), CAST(1 AS BIGINT), CAST(2 AS BIGINT), CAST(0 AS BIGINT), CAST(1.5 AS DOUBLE));\n"
    sql += "SELECT iterativelength(0, CAST(10 AS BIGINT), CAST(1 AS BIGINT));\n"  # no such overload
    r = run_sql(sql)
    assert r[0][0] == "result"
    assert r[1] == ("result", ["cheapest_path_length", "cheapest_path_length_double", "create_csr_edge",
                               "create_csr_vertex", "delete_csr", "iterativelength",
                               "reachability", "shortestpath", "shortestpath_with_edges"])
    assert r[2][0] == "error"
    if not torch.cuda.is_available():
        assert "no CPU fallback" in r[2][1] or "CUDA" in r[2][1]
    else:
        assert "does not exist" in r[2][1]
    for k in (3, 4, 5):  # on a GPU box the weighted create_csr_edge simply succeeds
        assert r[k][0] == "result" or ("No function matches" not in r[k][1] and "Binder" not in r[k][1]), r[k]
    assert r[6][0] == "error" and ("No function matches" in r[6][1] or "Binder" in r[6][1]), r[6]


def _write_csv(path, cols):
    with open(path, "w") as f:
        for row in zip(*cols):
            f.write(",".join("" if x is None else str(int(x)) for x in row) + "\n")


def _parse_list(s):
    s = s.strip()
    if s == "NULL":
        return None
    return [int(x) for x in s.strip("[]").split(",") if x.strip()]


@pytest.mark.gpu
@needs_host
def test_sql_pgq_functions_match_oracle(tmp_path):
    V, E = 2500, 40000
    src, dst = orc.gen_knows(77, V, E)
    v, e, eids = orc.csr_build(V, src, dst)
    n = 2300  # > 2 chunks of 1024 rows
    ps, pd = orc.gen_pairs(5, V, v, n)
    ps = ps.astype(object)
    pd = pd.astype(object)
    for i in range(0, n, 97):
        ps[i] = None  # NULL src
    for i in range(5, n, 131):
        pd[i] = ps[i]  # src == dst
    sv = np.array([x is not None for x in ps], dtype=np.uint8)
    ps_i = np.array([0 if x is None else x for x in ps], dtype=np.int64)
    pd_i = np.array([0 if x is None else x for x in pd], dtype=np.int64)
    _write_csv(tmp_path / "person.csv", [np.arange(V) * 7 + 3])  # sparse ids: dense id = rowid
    _write_csv(tmp_path / "knows.csv", [src * 7 + 3, dst * 7 + 3])
    _write_csv(tmp_path / "pairs.csv", [np.arange(n), ps, pd])
    sql = """LOAD '%s';
PRAGMA threads=1;
CREATE TABLE person(id BIGINT);
CREATE TABLE knows(p1 BIGINT, p2 BIGINT);
CREATE TABLE pairs(i BIGINT, s BIGINT, d BIGINT);
COPY person FROM '%s';
COPY knows FROM '%s';
COPY pairs FROM '%s';
SELECT sum(create_csr_edge(0, (SELECT count(*) FROM person),
    CAST((SELECT sum(create_csr_vertex(0, (SELECT count(*) FROM person), sub.dense_id, sub.cnt))
          FROM (SELECT p.rowid AS dense_id, count(k.p1) AS cnt FROM person p LEFT JOIN knows k ON k.p1 = p.id
                GROUP BY p.rowid) sub) AS BIGINT),
    a.rowid, b.rowid, k.rowid))
  FROM knows k JOIN person a ON k.p1 = a.id JOIN person b ON k.p2 = b.id;
SELECT i, iterativelength(0, (SELECT count(*) FROM person), s, d) FROM pairs ORDER BY i;
SELECT i, reachability(0, (SELECT count(*) FROM person), s, d) FROM pairs ORDER BY i;
SELECT i, shortestpath(0, (SELECT count(*) FROM person), s, d) FROM pairs ORDER BY i;
SELECT i, iterativelength(0, (SELECT count(*) FROM person), s, d, CAST(2 AS BIGINT), CAST(3 AS BIGINT)) FROM pairs ORDER BY i;
SELECT delete_csr(0);
SELECT iterativelength(0, CAST(5 AS BIGINT), CAST(1 AS BIGINT), CAST(2 AS BIGINT));
""" % (EXT, tmp_path / "person.csv", tmp_path / "knows.csv", tmp_path / "pairs.csv")
    r = run_sql(sql)
    kinds = [k for k, _ in r]
    assert kinds[:8] == ["result"] * 8, r[:8]
    assert r[8] == ("result", [str(E)])  # every edge row accepted
    eo, ev = orc.iterativelength(V, v, e, ps_i, pd_i, src_valid=sv)
    got = [row.split(",", 1)[1] for row in r[9][1]]
    assert got == ["NULL" if not ev[i] else str(int(eo[i])) for i in range(n)]
    er, erv = orc.reachability(V, v, e, ps_i, pd_i, src_valid=sv)
    got = [row.split(",", 1)[1] for row in r[10][1]]
    assert got == ["NULL" if not erv[i] else ("True" if er[i] else "False") for i in range(n)]
    off, ln, pv, vals = orc.shortestpath(V, v, e, eids, ps_i, pd_i, src_valid=sv, mode=0)
    got = [_parse_list(row.split(",", 1)[1]) for row in r[11][1]]
    exp = [None if not pv[i] else [int(x) for x in vals[int(off[i]):int(off[i]) + int(ln[i])]] for i in range(n)]
    assert got == exp
    bo, bv = orc.iterativelength_bounded(V, v, e, ps_i, pd_i, 2, 3, src_valid=sv)  # 6-argument form (R11)
    got = [row.split(",", 1)[1] for row in r[12][1]]
    assert got == ["NULL" if not bv[i] else str(int(bo[i])) for i in range(n)]
    assert r[13] == ("result", ["True"])
    assert r[14][0] == "error" and "does not exist" in r[14][1]


@pytest.mark.gpu
@needs_host
def test_sql_edge_ids_in_scan_order(tmp_path):
    """Direct scan (no join), threads=1: rows reach create_csr_edge in table order, so the
    edge-id-interleaved paths equal the oracle's byte for byte (R3, R7, R8)."""
    scale = 10
    V = 1 << scale
    src, dst = orc.gen_rmat(0x5eed0000 + scale, scale)
    E = len(src)
    v, e, eids = orc.csr_build(V, src, dst)
    n = 600
    ps, pd = orc.gen_pairs(9, V, v, n)
    _write_csv(tmp_path / "e.csv", [src, dst])
    _write_csv(tmp_path / "pairs.csv", [np.arange(n), ps, pd])
    sql = """LOAD '%s';
PRAGMA threads=1;
CREATE TABLE e(s BIGINT, d BIGINT);
CREATE TABLE pairs(i BIGINT, s BIGINT, d BIGINT);
COPY e FROM '%s';
COPY pairs FROM '%s';
SELECT sum(create_csr_edge(3, CAST(%d AS BIGINT), CAST(%d AS BIGINT), s, d, rowid)) FROM e;
SELECT i, shortestpath_with_edges(3, CAST(%d AS BIGINT), s, d) FROM pairs ORDER BY i;
""" % (EXT, tmp_path / "e.csv", tmp_path / "pairs.csv", V, E, V)
    r = run_sql(sql)
    assert [k for k, _ in r] == ["result"] * 8, r
    assert r[6] == ("result", [str(E)])
    off, ln, pv, vals = orc.shortestpath(V, v, e, eids, ps, pd, mode=1)
    got = [_parse_list(row.split(",", 1)[1]) for row in r[7][1]]
    exp = [None if not pv[i] else [int(x) for x in vals[int(off[i]):int(off[i]) + int(ln[i])]] for i in range(n)]
    assert got == exp


@pytest.mark.gpu
@needs_host
def test_sql_cheapest_path_length(tmp_path):
    """7-argument create_csr_edge (edge rowid + weight) and cheapest_path_length through SQL, BIGINT and
    DOUBLE weights, against the oracle's Bellman-Ford (rule R10).  The distance does not depend on
    the neighbour order, so no ORDER BY / threads=1 subtleties apply to the values."""
    scale = 10
    V = 1 << scale
    src, dst = orc.gen_rmat(0x5eed0000 + scale, scale)
    E = len(src)
    v, e, _ = orc.csr_build(V, src, dst, want_eids=False)
    wi = np.random.default_rng(21).integers(1, 500, size=E).astype(np.int64)
    n = 300
    ps, pd = orc.gen_pairs(10, V, v, n)
    ps = ps.astype(object)
    for i in range(0, n, 37):
        ps[i] = None
    sv = np.array([x is not None for x in ps], dtype=np.uint8)
    ps_i = np.array([0 if x is None else x for x in ps], dtype=np.int64)
    _write_csv(tmp_path / "e.csv", [src, dst, wi])
    _write_csv(tmp_path / "pairs.csv", [np.arange(n), ps, pd])
    sql = """LOAD '%s';
PRAGMA threads=1;
CREATE TABLE e(s BIGINT, d BIGINT, w BIGINT);
CREATE TABLE pairs(i BIGINT, s BIGINT, d BIGINT);
COPY e FROM '%s';
COPY pairs FROM '%s';
SELECT sum(create_csr_edge(4, CAST(%d AS BIGINT), CAST(%d AS BIGINT), s, d, rowid, w)) FROM e;
SELECT i, cheapest_path_length(4, CAST(%d AS BIGINT), s, d) FROM pairs ORDER BY i;
SELECT cheapest_path_length_double(4, CAST(%d AS BIGINT), s, d) FROM pairs LIMIT 1;
SELECT sum(create_csr_edge(5, CAST(%d AS BIGINT), CAST(%d AS BIGINT), s, d, rowid, CAST(w AS DOUBLE) / 4)) FROM e;
SELECT i, cheapest_path_length_double(5, CAST(%d AS BIGINT), s, d) FROM pairs ORDER BY i;
SELECT iterativelength(5, CAST(%d AS BIGINT), s, d) FROM pairs WHERE i = 1;
""" % (EXT, tmp_path / "e.csv", tmp_path / "pairs.csv", V, E, V, V, V, E, V, V)
    r = run_sql(sql)
    assert [k for k, _ in r[:8]] == ["result"] * 8, r[:8]
    assert r[6] == ("result", [str(E)])
    wc = orc.csr_weights(V, v, src, wi)
    eo, ev, _ = orc.cheapest_path_length(V, v, e, wc, ps_i, pd, src_valid=sv)
    got = [row.split(",", 1)[1] for row in r[7][1]]
    assert got == ["NULL" if not ev[i] else str(int(eo[i])) for i in range(n)]
    assert r[8][0] == "error" and "BIGINT weights" in r[8][1]
    assert r[9] == ("result", [str(E)])
    wcd = orc.csr_weights(V, v, src, wi.astype(np.float64) / 4)
    do, dv, _ = orc.cheapest_path_length(V, v, e, wcd, ps_i, pd, src_valid=sv)
    got = [row.split(",", 1)[1] for row in r[10][1]]
    assert [g == "NULL" for g in got] == [not dv[i] for i in range(n)]
    assert [float(g) for i, g in enumerate(got) if dv[i]] == [float(do[i]) for i in range(n) if dv[i]]
    assert r[11][0] == "result"  # the unweighted functions work on a weighted CSR too


@pytest.mark.gpu
@needs_host
def test_sql_functions_against_the_engines_own_recursive_cte(tmp_path):
    """Inside ONE reference-DuckDB session: the recursive-CTE BFS of /root/reference/benchmark/ldbc/queries/
    bi-10-shortestpath.sql:8-31 (the only BFS the mounted reference can run) joined against iterativelength /
    reachability / the upper-bounded iterativelength evaluated by the extension on the same pairs -- the engine
    itself counts the disagreements."""
    V, E = 900, 5400
    src, dst = orc.gen_knows(99, V, E)
    v, e, _ = orc.csr_build(V, src, dst)
    n = 1500
    ps, pd = orc.gen_pairs(12, V, v, n)
    ps[:] = ps[:20][np.arange(n) % 20]  # 20 start vertices keep the CTE small
    _write_csv(tmp_path / "person.csv", [np.arange(V)])  # id = dense row id
    _write_csv(tmp_path / "knows.csv", [src, dst])
    _write_csv(tmp_path / "pairs.csv", [ps, pd])
    cte = """WITH RECURSIVE friends(startPerson, hopCount, friend) AS (
    SELECT s, 0, s FROM (SELECT DISTINCT s FROM pairs) t
  UNION
    SELECT f.startPerson, f.hopCount+1, k.p2 FROM friends f, knows k WHERE f.friend = k.p1 AND f.hopCount < %d
), fs AS (SELECT startPerson, min(hopCount) AS hopCount, friend FROM friends WHERE hopCount > 0
          GROUP BY startPerson, friend),
ours AS (SELECT s, d, %s AS got FROM pairs)
SELECT count(*), sum(CASE WHEN %s THEN 0 ELSE 1 END)
  FROM ours o LEFT JOIN fs ON fs.startPerson = o.s AND fs.friend = o.d;
"""
    same_len = "coalesce(CASE WHEN o.s = o.d THEN 0 ELSE fs.hopCount END, -1) = coalesce(o.got, -1)"
    same_reach = "(o.s = o.d OR fs.hopCount IS NOT NULL) = o.got"
    vs = "CAST(%d AS BIGINT)" % V
    sql = """LOAD '%s';
PRAGMA threads=1;
CREATE TABLE person(id BIGINT);
CREATE TABLE knows(p1 BIGINT, p2 BIGINT);
CREATE TABLE pairs(s BIGINT, d BIGINT);
COPY person FROM '%s';
COPY knows FROM '%s';
COPY pairs FROM '%s';
SELECT sum(create_csr_edge(0, (SELECT count(*) FROM person),
    CAST((SELECT sum(create_csr_vertex(0, (SELECT count(*) FROM person), sub.dense_id, sub.cnt))
          FROM (SELECT p.rowid AS dense_id, count(k.p1) AS cnt FROM person p LEFT JOIN knows k ON k.p1 = p.id
                GROUP BY p.rowid) sub) AS BIGINT),
    a.rowid, b.rowid, k.rowid))
  FROM knows k JOIN person a ON k.p1 = a.id JOIN person b ON k.p2 = b.id;
""" % (EXT, tmp_path / "person.csv", tmp_path / "knows.csv", tmp_path / "pairs.csv")
    sql += cte % (24, "iterativelength(0, %s, s, d)" % vs, same_len)
    sql += cte % (24, "reachability(0, %s, s, d)" % vs, same_reach)
    for upper in (1, 2, 3):
        sql += cte % (upper, "iterativelength(0, %s, s, d, CAST(0 AS BIGINT), CAST(%d AS BIGINT))" % (vs, upper),
                      same_len)
    sql += "SELECT delete_csr(0);\n"
    r = run_sql(sql)
    assert all(k == "result" for k, _ in r), [x for x in r if x[0] != "result"][:2]
    assert r[8] == ("result", [str(E)])
    for k in range(9, 14):
        assert r[k] == ("result", ["%d,0" % n]), (k, r[k])


@needs_host
def test_oracle_against_live_recursive_cte(tmp_path):
    """CPU: the same comparison with the ORACLE's answers loaded as a table -- the reference DuckDB built from
    /root/reference runs its recursive-CTE BFS live and counts the disagreements with the oracle's
    iterativelength / reachability / upper-bounded iterativelength (pins the oracle beyond the committed
    fixtures; skipped where the reference host is not built)."""
    V, E = 900, 5400
    src, dst = orc.gen_knows(99, V, E)
    v, e, _ = orc.csr_build(V, src, dst)
    n = 1500
    ps, pd = orc.gen_pairs(12, V, v, n)
    ps[:] = ps[:20][np.arange(n) % 20]
    _write_csv(tmp_path / "knows.csv", [src, dst])
    _write_csv(tmp_path / "pairs.csv", [ps, pd])
    cte = """WITH RECURSIVE friends(startPerson, hopCount, friend) AS (
    SELECT s, 0, s FROM (SELECT DISTINCT s FROM pairs) t
  UNION
    SELECT f.startPerson, f.hopCount+1, k.p2 FROM friends f, knows k WHERE f.friend = k.p1 AND f.hopCount < %d
), fs AS (SELECT startPerson, min(hopCount) AS hopCount, friend FROM friends WHERE hopCount > 0
          GROUP BY startPerson, friend)
SELECT count(*), sum(CASE WHEN %s THEN 0 ELSE 1 END)
  FROM %s o LEFT JOIN fs ON fs.startPerson = o.s AND fs.friend = o.d;
"""
    same_len = "coalesce(CASE WHEN o.s = o.d THEN 0 ELSE fs.hopCount END, -1) = coalesce(o.got, -1)"
    same_reach = "(o.s = o.d OR fs.hopCount IS NOT NULL) = (o.got = 1)"
    sql = "CREATE TABLE knows(p1 BIGINT, p2 BIGINT);\nCOPY knows FROM '%s';\n" % (tmp_path / "knows.csv")
    sql += "CREATE TABLE pairs(s BIGINT, d BIGINT);\nCOPY pairs FROM '%s';\n" % (tmp_path / "pairs.csv")
    nq = 0

    def table(name, vals):
        _write_csv(tmp_path / (name + ".csv"), [ps, pd, vals])
        return "CREATE TABLE %s(s BIGINT, d BIGINT, got BIGINT);\nCOPY %s FROM '%s';\n" % (
            name, name, tmp_path / (name + ".csv"))

    full, fv = orc.iterativelength(V, v, e, ps, pd)
    sql += table("t_len", [None if not fv[i] else full[i] for i in range(n)]) + cte % (24, same_len, "t_len")
    reach, _ = orc.reachability(V, v, e, ps, pd)
    sql += table("t_reach", reach) + cte % (24, same_reach, "t_reach")
    nq += 2
    for upper in (1, 2, 3):
        bo, bv = orc.iterativelength_bounded(V, v, e, ps, pd, 0, upper)
        name = "t_b%d" % upper
        sql += table(name, [None if not bv[i] else bo[i] for i in range(n)]) + cte % (upper, same_len, name)
        nq += 1
    r = run_sql(sql)
    assert all(k == "result" for k, _ in r), [x for x in r if x[0] != "result"][:2]
    verdicts = [rows for _, rows in r if rows and "," in rows[0]]
    assert verdicts == [["%d,0" % n]] * nq, verdicts


@needs_host
def test_oracle_csr_against_engine_sql(tmp_path):
    """CPU: rule R3 (neighbour order inside a vertex = arrival order of the rows) restated in SQL and run by the
    reference DuckDB: `e` = targets ordered by (source, rowid), `v` = running count of the sources."""
    V, E = 300, 4000
    rng = np.random.default_rng(8)
    src = rng.integers(0, V, E)
    dst = rng.integers(0, V, E)
    src[::7] = 5  # a long neighbour list with many duplicates
    v, e, _ = orc.csr_build(V, src, dst)
    _write_csv(tmp_path / "knows.csv", [src, dst])
    sql = "PRAGMA threads=1;\nCREATE TABLE knows(p1 BIGINT, p2 BIGINT);\nCOPY knows FROM '%s';\n" % (tmp_path / "knows.csv")
    sql += "SELECT p2 FROM knows ORDER BY p1, rowid;\n"
    sql += "SELECT p1, count(*) FROM knows GROUP BY p1 ORDER BY p1;\n"
    r = run_sql(sql)
    assert all(k == "result" for k, _ in r), r
    assert [int(x) for x in r[3][1]] == e.tolist()
    deg = np.zeros(V, dtype=np.int64)
    for row in r[4][1]:
        a, b = row.split(",")
        deg[int(a)] = int(b)
    expect_v = np.concatenate([[0], np.cumsum(deg), [E]])
    assert expect_v.tolist() == v.tolist()


# ---- round 2: the parent tie-break (R7), the edge-id choice (R8), the lower bound (R11) and the weighted
# distances (R10) restated in SQL and evaluated by the reference DuckDB ------------------------------------
def _path_walk_sql(lmax):
    """SQL text (no PGQ function involved) that reconstructs, for every row of pairs(i, s, d), the path rule R7
    describes: dist = the recursive-CTE BFS of benchmark/ldbc/queries/bi-10-shortestpath.sql:8-31; then from the
    destination backwards  node[k-1] = min(u : dist(s,u) = k-1 and an edge u -> node[k] exists)  ("the smallest
    vertex id of the previous level's frontier that has an edge to n"), one INSERT per level; and R8: the edge
    of a step = min(rowid) over the parallel edges node[k] -> node[k+1] (lowest CSR offset = earliest row)."""
    sql = """CREATE TABLE dist AS SELECT startPerson AS s, friend AS v, min(hopCount) AS h FROM (
WITH RECURSIVE friends(startPerson, hopCount, friend) AS (
    SELECT s, 0, s FROM (SELECT DISTINCT s FROM pairs) t
  UNION
    SELECT f.startPerson, f.hopCount+1, k.p2 FROM friends f, knows k WHERE f.friend = k.p1 AND f.hopCount < %d
) SELECT * FROM friends) x GROUP BY startPerson, friend;
CREATE TABLE walk AS SELECT p.i AS i, p.s AS s, dd.h AS k, p.d AS node FROM pairs p JOIN dist dd ON dd.s = p.s AND dd.v = p.d WHERE p.s <> p.d;
""" % lmax
    for k in range(lmax, 0, -1):
        sql += ("INSERT INTO walk SELECT w.i, w.s, w.k - 1, min(k2.p1) FROM walk w JOIN knows k2 ON k2.p2 = w.node "
                "JOIN dist d2 ON d2.s = w.s AND d2.v = k2.p1 AND d2.h = w.k - 1 WHERE w.k = %d GROUP BY w.i, w.s, w.k;\n" % k)
    sql += "SELECT i, k, node FROM walk ORDER BY i, k;\n"
    sql += ("SELECT w.i, w.k, min(k3.rowid) FROM walk w JOIN walk w2 ON w2.i = w.i AND w2.k = w.k + 1 "
            "JOIN knows k3 ON k3.p1 = w.node AND k3.p2 = w2.node GROUP BY w.i, w.k ORDER BY w.i, w.k;\n")
    return sql


def _collect_walk(rows_nodes, rows_edges, n):
    nodes = [dict() for _ in range(n)]
    edges = [dict() for _ in range(n)]
    for row in rows_nodes:
        i, k, node = (int(x) for x in row.split(","))
        nodes[i][k] = node
    for row in rows_edges:
        i, k, eid = (int(x) for x in row.split(","))
        edges[i][k] = eid
    out0, out1 = [], []
    for i in range(n):
        if not nodes[i]:
            out0.append(None)
            out1.append(None)
            continue
        L = max(nodes[i])
        assert sorted(nodes[i]) == list(range(L + 1)) and sorted(edges[i]) == list(range(L)), i
        out0.append([nodes[i][k] for k in range(L + 1)])
        inter = []
        for k in range(L):
            inter += [nodes[i][k], edges[i][k]]
        out1.append(inter + [nodes[i][L]])
    return out0, out1


def _tie_break_graph():
    """knows-shaped graph with many equal-length alternatives and parallel edges, so R7 / R8 actually decide."""
    V, E = 400, 3600
    src, dst = orc.gen_knows(1234, V, E)
    rng = np.random.default_rng(4)
    dup = rng.integers(0, E, 500)  # parallel edges: repeat 500 rows somewhere else in the table
    pos = rng.integers(0, E, 500)
    src = np.insert(src, pos, src[dup])
    dst = np.insert(dst, pos, dst[dup])
    return V, src, dst


@needs_host
def test_oracle_parent_rule_and_edge_ids_against_engine_sql(tmp_path):
    """CPU: R7 (parent = smallest-id vertex of the previous frontier with an edge to n) and R8 (edge of a step =
    lowest CSR offset among parallel edges = earliest row) restated in plain SQL, run by the reference DuckDB,
    against the oracle's shortestpath, vertex-only and edge-interleaved, list by list."""
    V, src, dst = _tie_break_graph()
    v, e, eids = orc.csr_build(V, src, dst)
    n = 600
    ps, pd = orc.gen_pairs(31, V, v, n)
    ps[:] = ps[:25][np.arange(n) % 25]
    _write_csv(tmp_path / "knows.csv", [src, dst])
    _write_csv(tmp_path / "pairs.csv", [np.arange(n), ps, pd])
    sql = "PRAGMA threads=1;\nCREATE TABLE knows(p1 BIGINT, p2 BIGINT);\nCOPY knows FROM '%s';\n" % (tmp_path / "knows.csv")
    sql += "CREATE TABLE pairs(i BIGINT, s BIGINT, d BIGINT);\nCOPY pairs FROM '%s';\n" % (tmp_path / "pairs.csv")
    lmax = 12
    sql += _path_walk_sql(lmax)
    r = run_sql(sql)
    assert all(k == "result" for k, _ in r), [x for x in r if x[0] != "result"][:2]
    sql0, sql1 = _collect_walk(r[-2][1], r[-1][1], n)
    for mode, want in ((0, sql0), (1, sql1)):
        off, ln, pv, vals = orc.shortestpath(V, v, e, eids, ps, pd, mode=mode)
        got = [None if not pv[i] else [int(x) for x in vals[int(off[i]):int(off[i]) + int(ln[i])]] for i in range(n)]
        ties = 0
        for i in range(n):
            if ps[i] == pd[i]:
                assert got[i] == [int(ps[i])]
                continue
            assert got[i] == want[i], (mode, i, got[i], want[i])
            ties += got[i] is not None and len(got[i]) > 2
        assert ties > 100  # the rule was exercised on paths with intermediate vertices
    lens = [len(p) - 1 for p in sql0 if p]
    assert max(lens) < lmax


@needs_host
def test_oracle_lower_bound_and_weights_against_engine_sql(tmp_path):
    """CPU: R11's lower bound (a pair whose BFS length is below `lower` gives NULL) against the recursive CTE,
    and R10 (cheapest_path_length = Bellman-Ford fixed point over BIGINT weights) against relaxation rounds
    written as plain SQL (one CREATE TABLE ... GROUP BY per round until nothing changes)."""
    V, E = 300, 2400
    src, dst = orc.gen_knows(777, V, E)
    v, e, _ = orc.csr_build(V, src, dst, want_eids=False)
    w = np.random.default_rng(3).integers(1, 40, size=E).astype(np.int64)
    n = 500
    ps, pd = orc.gen_pairs(17, V, v, n)
    ps[:] = ps[:10][np.arange(n) % 10]
    _write_csv(tmp_path / "knows.csv", [src, dst, w])
    _write_csv(tmp_path / "pairs.csv", [np.arange(n), ps, pd])
    sql = "PRAGMA threads=1;\nCREATE TABLE knows(p1 BIGINT, p2 BIGINT, w BIGINT);\nCOPY knows FROM '%s';\n" % (tmp_path / "knows.csv")
    sql += "CREATE TABLE pairs(i BIGINT, s BIGINT, d BIGINT);\nCOPY pairs FROM '%s';\n" % (tmp_path / "pairs.csv")
    lower, upper = 3, 4
    sql += """WITH RECURSIVE friends(startPerson, hopCount, friend) AS (
    SELECT s, 0, s FROM (SELECT DISTINCT s FROM pairs) t
  UNION
    SELECT f.startPerson, f.hopCount+1, k.p2 FROM friends f, knows k WHERE f.friend = k.p1 AND f.hopCount < %d
), fs AS (SELECT startPerson, min(hopCount) AS h, friend FROM friends WHERE hopCount > 0 GROUP BY startPerson, friend)
SELECT p.i, CASE WHEN p.s = p.d THEN (CASE WHEN %d <= 0 THEN 0 ELSE NULL END) WHEN fs.h >= %d THEN fs.h ELSE NULL END
  FROM pairs p LEFT JOIN fs ON fs.startPerson = p.s AND fs.friend = p.d ORDER BY p.i;
""" % (upper, lower, lower)
    rounds = 40
    sql += "CREATE TABLE d0 AS SELECT DISTINCT s, s AS v, CAST(0 AS BIGINT) AS c FROM pairs;\n"
    for k in range(rounds):
        sql += ("CREATE TABLE d%d AS SELECT s, v, min(c) AS c FROM (SELECT s, v, c FROM d%d UNION ALL "
                "SELECT d.s, k.p2, d.c + k.w FROM d%d d JOIN knows k ON k.p1 = d.v) x GROUP BY s, v;\n" % (k + 1, k, k))
    sql += "SELECT (SELECT sum(c) FROM d%d), (SELECT sum(c) FROM d%d), (SELECT count(*) FROM d%d), (SELECT count(*) FROM d%d);\n" % (
        rounds, rounds - 1, rounds, rounds - 1)
    sql += "SELECT p.i, d.c FROM pairs p LEFT JOIN d%d d ON d.s = p.s AND d.v = p.d ORDER BY p.i;\n" % rounds
    r = run_sql(sql)
    assert all(k == "result" for k, _ in r), [x for x in r if x[0] != "result"][:2]
    bo, bv = orc.iterativelength_bounded(V, v, e, ps, pd, lower, upper)
    got = [row.split(",", 1)[1] for row in r[5][1]]
    assert got == ["NULL" if not bv[i] else str(int(bo[i])) for i in range(n)]
    assert any(g != "NULL" for g in got) and any(bv[i] == 0 and orc.iterativelength(V, v, e, ps[i:i + 1], pd[i:i + 1])[1][0]
                                                 for i in range(0, n, 7))  # some pairs fell below the lower bound
    a, b, c, d = r[-2][1][0].split(",")
    assert a == b and c == d, "relaxation did not reach its fixed point in %d rounds" % rounds
    wc = orc.csr_weights(V, v, src, w)
    eo, ev, _ = orc.cheapest_path_length(V, v, e, wc, ps, pd)
    got = [row.split(",", 1)[1] for row in r[-1][1]]
    assert got == ["NULL" if not ev[i] else str(int(eo[i])) for i in range(n)]


@pytest.mark.gpu
@needs_host
def test_sql_paths_against_the_engines_own_walk(tmp_path):
    """GPU, ONE reference-DuckDB session: shortestpath / shortestpath_with_edges evaluated by the extension next to
    the plain-SQL restatement of R7 / R8 (see _path_walk_sql) on the same tables; plus the lower bound (R11) and
    cheapest_path_length (R10) against their SQL restatements."""
    V, src, dst = _tie_break_graph()
    E = len(src)
    n = 600
    v, e, eids = orc.csr_build(V, src, dst)
    ps, pd = orc.gen_pairs(31, V, v, n)
    ps[:] = ps[:25][np.arange(n) % 25]
    w = np.random.default_rng(3).integers(1, 40, size=E).astype(np.int64)
    _write_csv(tmp_path / "knows.csv", [src, dst, w])
    _write_csv(tmp_path / "pairs.csv", [np.arange(n), ps, pd])
    vs = "CAST(%d AS BIGINT)" % V
    sql = "LOAD '%s';\nPRAGMA threads=1;\n" % EXT
    sql += "CREATE TABLE knows(p1 BIGINT, p2 BIGINT, w BIGINT);\nCOPY knows FROM '%s';\n" % (tmp_path / "knows.csv")
    sql += "CREATE TABLE pairs(i BIGINT, s BIGINT, d BIGINT);\nCOPY pairs FROM '%s';\n" % (tmp_path / "pairs.csv")
    sql += "SELECT sum(create_csr_edge(6, %s, CAST(%d AS BIGINT), p1, p2, rowid, w)) FROM knows;\n" % (vs, E)
    sql += "SELECT i, shortestpath(6, %s, s, d) FROM pairs ORDER BY i;\n" % vs
    sql += "SELECT i, shortestpath_with_edges(6, %s, s, d) FROM pairs ORDER BY i;\n" % vs
    sql += "SELECT i, iterativelength(6, %s, s, d, CAST(3 AS BIGINT), CAST(4 AS BIGINT)) FROM pairs ORDER BY i;\n" % vs
    sql += "SELECT i, cheapest_path_length(6, %s, s, d) FROM pairs ORDER BY i;\n" % vs
    base = 7  # statements so far (results r[0..base+3])
    sql += _path_walk_sql(12)
    sql += """WITH fs AS (SELECT s AS startPerson, v AS friend, h FROM dist WHERE h > 0 AND h <= 4)
SELECT p.i, CASE WHEN p.s = p.d THEN NULL WHEN fs.h >= 3 THEN fs.h ELSE NULL END
  FROM pairs p LEFT JOIN fs ON fs.startPerson = p.s AND fs.friend = p.d ORDER BY p.i;
"""
    rounds = 40
    sql += "CREATE TABLE d0 AS SELECT DISTINCT s, s AS v, CAST(0 AS BIGINT) AS c FROM pairs;\n"
    for k in range(rounds):
        sql += ("CREATE TABLE d%d AS SELECT s, v, min(c) AS c FROM (SELECT s, v, c FROM d%d UNION ALL "
                "SELECT d.s, k.p2, d.c + k.w FROM d%d d JOIN knows k ON k.p1 = d.v) x GROUP BY s, v;\n" % (k + 1, k, k))
    sql += "SELECT (SELECT sum(c) FROM d%d) = (SELECT sum(c) FROM d%d);\n" % (rounds, rounds - 1)
    sql += "SELECT p.i, d.c FROM pairs p LEFT JOIN d%d d ON d.s = p.s AND d.v = p.d ORDER BY p.i;\n" % rounds
    r = run_sql(sql)
    assert all(k == "result" for k, _ in r), [x for x in r if x[0] != "result"][:2]
    assert r[base - 1] == ("result", [str(E)])
    ext0 = [_parse_list(row.split(",", 1)[1]) for row in r[base][1]]
    ext1 = [_parse_list(row.split(",", 1)[1]) for row in r[base + 1][1]]
    ext_lb = [row.split(",", 1)[1] for row in r[base + 2][1]]
    ext_cp = [row.split(",", 1)[1] for row in r[base + 3][1]]
    walk_edges = [i for i, (k, rows) in enumerate(r) if k == "result" and rows and rows[0].count(",") == 2]
    sql0, sql1 = _collect_walk(r[walk_edges[0]][1], r[walk_edges[1]][1], n)
    for i in range(n):
        if ps[i] == pd[i]:
            assert ext0[i] == [int(ps[i])] and ext1[i] == [int(ps[i])]
        else:
            assert ext0[i] == sql0[i], (i, ext0[i], sql0[i])
            assert ext1[i] == sql1[i], (i, ext1[i], sql1[i])
    lb_sql = [row.split(",", 1)[1] for row in r[walk_edges[1] + 1][1]]
    assert ext_lb == lb_sql
    assert r[-2] == ("result", ["True"])
    cp_sql = [row.split(",", 1)[1] for row in r[-1][1]]
    assert ext_cp == cp_sql


# ---- round 2: the shim under concurrent entry, restartable builds, edge-less graphs, constant inputs ---------
def _knows_setup_sql(tmp_path, V, src, dst, ps, pd, threads):
    _write_csv(tmp_path / "person.csv", [np.arange(V)])
    _write_csv(tmp_path / "knows.csv", [src, dst])
    _write_csv(tmp_path / "pairs.csv", [np.arange(len(ps)), ps, pd])
    return """LOAD '%s';
PRAGMA threads=%d;
CREATE TABLE person(id BIGINT);
CREATE TABLE knows(p1 BIGINT, p2 BIGINT);
CREATE TABLE pairs(i BIGINT, s BIGINT, d BIGINT);
COPY person FROM '%s';
COPY knows FROM '%s';
COPY pairs FROM '%s';
""" % (EXT, threads, tmp_path / "person.csv", tmp_path / "knows.csv", tmp_path / "pairs.csv")


@pytest.mark.gpu
@needs_host
@pytest.mark.parametrize("verify", [False, True])
def test_sql_concurrent_entry(tmp_path, verify):
    """PRAGMA threads=4 (and the reference CI's PRAGMA verify_parallelism, src/function/pragma/pragma_functions.cpp:
    305,311): the query functions are entered by several workers at once, their rows meet in the slot's batcher and
    every row still gets its own answer.  The CSR is built with threads=1 first (its neighbour order is the row
    arrival order), and lengths / reachability do not depend on that order anyway; paths are compared for
    threads=4 too because the CSR was built single-threaded."""
    V, E = 3000, 60000
    src, dst = orc.gen_knows(4711, V, E)
    v, e, eids = orc.csr_build(V, src, dst)
    n = 20000  # ~20 chunks of 1024 rows
    ps, pd = orc.gen_pairs(13, V, v, n)
    ps[::53] = pd[::53]
    vs = "CAST(%d AS BIGINT)" % V
    sql = _knows_setup_sql(tmp_path, V, src, dst, ps, pd, 1)
    sql += "SELECT sum(create_csr_edge(0, %s, CAST(%d AS BIGINT), p1, p2, rowid)) FROM knows;\n" % (vs, E)
    sql += "PRAGMA threads=4;\n"
    if verify:
        sql += "PRAGMA verify_parallelism;\n"
    sql += "SELECT i, iterativelength(0, %s, s, d) FROM pairs ORDER BY i;\n" % vs
    sql += "SELECT i, reachability(0, %s, s, d) FROM pairs ORDER BY i;\n" % vs
    sql += "SELECT i, shortestpath_with_edges(0, %s, s, d) FROM pairs WHERE i < 6000 ORDER BY i;\n" % vs
    sql += "SELECT sum(iterativelength(0, %s, s, d)), count(*) FROM pairs;\n" % vs
    sql += "SELECT pgq_batcher_stats(0);\n"
    r = run_sql(sql)
    assert all(k == "result" for k, _ in r), [x for x in r if x[0] != "result"][:2]
    res = [rows for k, rows in r if rows and rows[0].count(",") >= 1]
    eo, ev = orc.iterativelength(V, v, e, ps, pd)
    got = [row.split(",", 1)[1] for row in res[0]]
    assert got == ["NULL" if not ev[i] else str(int(eo[i])) for i in range(n)]
    got = [row.split(",", 1)[1] for row in res[1]]
    assert got == ["True" if ev[i] else "False" for i in range(n)]
    off, ln, pv, vals = orc.shortestpath(V, v, e, eids, ps[:6000], pd[:6000], mode=1)
    got = [_parse_list(row.split(",", 1)[1]) for row in res[2]]
    exp = [None if not pv[i] else [int(x) for x in vals[int(off[i]):int(off[i]) + int(ln[i])]] for i in range(6000)]
    assert got == exp
    assert res[3] == ["%d,%d" % (int(eo[ev == 1].sum()), n)]
    stats = r[-1][1][0]
    calls, gpu_calls = (int(x.split("=")[1]) for x in stats.split())
    assert calls >= 3 * (n // 1024) and gpu_calls <= calls


@pytest.mark.gpu
@needs_host
def test_sql_build_is_restartable_and_edgeless_graphs_answer(tmp_path):
    """(1) a create_csr_edge statement that aborts part-way (NULL row) must not wedge the slot -- the
    normal build statements simply run again; (2) a graph with vertices and ZERO edge rows finalizes (queries
    give NULL / false, not an internal error); (3) all-constant arguments give a constant result."""
    V = 50
    src = np.arange(V - 1)
    dst = src + 1
    _write_csv(tmp_path / "person.csv", [np.arange(V)])
    _write_csv(tmp_path / "knows.csv", [src, dst])
    bad_src = src.astype(object)
    bad_src[30] = None
    _write_csv(tmp_path / "bad.csv", [bad_src, dst])
    vs = "CAST(%d AS BIGINT)" % V
    build = ("SELECT sum(create_csr_edge(1, (SELECT count(*) FROM person), "
             "CAST((SELECT sum(create_csr_vertex(1, (SELECT count(*) FROM person), sub.dense_id, sub.cnt)) "
             "FROM (SELECT p.rowid AS dense_id, count(k.p1) AS cnt FROM person p LEFT JOIN %s k ON k.p1 = p.id "
             "GROUP BY p.rowid) sub) AS BIGINT), k.p1, k.p2)) FROM %s k;\n")
    sql = "LOAD '%s';\nPRAGMA threads=1;\n" % EXT
    sql += "CREATE TABLE person(id BIGINT);\nCOPY person FROM '%s';\n" % (tmp_path / "person.csv")
    sql += "CREATE TABLE knows(p1 BIGINT, p2 BIGINT);\nCOPY knows FROM '%s';\n" % (tmp_path / "knows.csv")
    sql += "CREATE TABLE bad(p1 BIGINT, p2 BIGINT);\nCOPY bad FROM '%s';\n" % (tmp_path / "bad.csv")
    sql += "CREATE TABLE nobody(p1 BIGINT, p2 BIGINT);\n"
    sql += build % ("bad", "bad")          # 9: aborts on the NULL row
    sql += build % ("knows", "knows")      # 9: the same statement on good rows just works
    sql += "SELECT iterativelength(1, %s, CAST(0 AS BIGINT), CAST(49 AS BIGINT));\n" % vs  # 10: constant args
    sql += build % ("knows", "knows")      # 11: rebuilding a finished CSR works too
    sql += "SELECT shortestpath(1, %s, CAST(3 AS BIGINT), CAST(6 AS BIGINT));\n" % vs  # 12
    # an edge-less graph: create_csr_vertex rows only (the edge query returns no row, so create_csr_edge never runs)
    sql += ("SELECT sum(create_csr_vertex(2, (SELECT count(*) FROM person), sub.dense_id, sub.cnt)) FROM "
            "(SELECT p.rowid AS dense_id, count(k.p1) AS cnt FROM person p LEFT JOIN nobody k ON k.p1 = p.id "
            "GROUP BY p.rowid) sub;\n")    # 13
    sql += "SELECT iterativelength(2, %s, id, id + 1), reachability(2, %s, id, id + 1), iterativelength(2, %s, id, id) FROM person WHERE id < 3 ORDER BY id;\n" % (vs, vs, vs)  # 14
    r = run_sql(sql)
    assert r[9][0] == "error" and "NULL" in r[9][1], r[9]
    assert r[10] == ("result", [str(V - 1)]), r[10]
    assert r[11] == ("result", ["49"]), r[11]
    assert r[12] == ("result", [str(V - 1)]), r[12]
    assert r[13] == ("result", ["[3, 4, 5, 6]"]), r[13]
    assert r[14] == ("result", ["0"]), r[14]
    assert r[15] == ("result", ["NULL,False,0"] * 3), r[15]
