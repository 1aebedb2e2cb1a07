"""the SQL/PGQ rewriter (duckdb-pgq.old_b200/ext/pgq_rewrite*.hpp) — CREATE PROPERTY GRAPH and
GRAPH_TABLE ( ... MATCH ... COLUMNS ... ) become the plain SQL that calls the path's scalar functions.

CPU: the rewriter on its own (pgq_rewrite_cli, no DuckDB, no GPU) and, where the reference DuckDB host was built, the
rewritten statements run by that engine with SQL macros standing in for the scalar functions (shape of the SQL only).
GPU: the same statements through the extension, answers against the oracle."""
import os
import subprocess

import numpy as np
import pytest

import orc
from test_sql_extension import DRIVER, EXT, needs_host, run_sql, _parse_list, _write_csv

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "duckdb-pgq.old_b200", "ext", "pgq_rewrite_cli")

GRAPH = ("CREATE PROPERTY GRAPH snb VERTEX TABLES (person KEY (id) LABEL Person) EDGE TABLES (knows SOURCE KEY (p1) "
         "REFERENCES person (id) DESTINATION KEY (p2) REFERENCES person (id) LABEL knows);\n")


def rewrite(sql):
    """-> list of ('passthrough', text) | ('rewritten', [statements]) | ('error', message)"""
    if not os.path.exists(CLI):
        subprocess.run(["make", "-s", "-C", os.path.dirname(CLI), "pgq_rewrite_cli"], check=True)
    out = subprocess.run([CLI], input=sql, capture_output=True, text=True, timeout=60)
    assert out.returncode == 0, out.stderr
    res, cur = [], None
    for line in out.stdout.splitlines():
        if line.startswith("-- passthrough "):
            cur = []
            res.append(("passthrough", cur))
        elif line.startswith("-- rewritten "):
            cur = []
            res.append(("rewritten", cur))
        elif line.startswith("-- error "):
            res.append(("error", line.split(":", 1)[1].strip()))
            cur = None
        elif line.startswith("-- done "):
            cur = None
        elif cur is not None:
            cur.append(line)
    return res


def test_create_property_graph_and_passthrough():
    r = rewrite(GRAPH + "SELECT 1;\nSELECT 'GRAPH_TABLE (' AS not_a_call;\nDROP PROPERTY GRAPH snb;\n"
                "DROP PROPERTY GRAPH snb;\nDROP PROPERTY GRAPH IF EXISTS snb;\n")
    assert r[0] == ("rewritten", [])  # catalog only: nothing for the engine to run
    assert r[1] == ("passthrough", ["SELECT 1;"])
    assert r[2][0] == "passthrough"  # the words inside a string literal are not syntax
    assert r[3] == ("rewritten", [])
    assert r[4][0] == "error" and "does not exist" in r[4][1]
    assert r[5] == ("rewritten", [])


def test_any_shortest_becomes_the_csr_template():
    q = ("SELECT * FROM GRAPH_TABLE (snb MATCH p = ANY SHORTEST (a:Person WHERE a.id = 3)-[k:knows]->*(b:Person) "
         "COLUMNS (a.id AS src, b.id AS dst, path_length(p) AS len, vertices(p) AS vs)) gt ORDER BY dst;\n")
    kind, st = rewrite(GRAPH + q)[1]
    assert kind == "rewritten" and len(st) == 2
    main = st[0]
    # the CTE builds the CSR with the reference's query template: dense ids are rowids, degrees from a LEFT JOIN
    assert main.startswith("WITH __pgq_csr1000 AS (SELECT create_csr_edge(1000, (SELECT count(*) FROM person), CAST((SELECT "
                           "sum(create_csr_vertex(1000, (SELECT count(*) FROM person), sub.dense_id, sub.cnt))")
    assert "s.rowid, d.rowid, e.rowid) AS temp FROM knows e JOIN person s ON s.id = e.p1 JOIN person d ON d.id = e.p2)" in main
    # the path functions cannot run before the CSR is complete: their source argument depends on the aggregate
    # (and |V| is a column of that one-row aggregate: a scalar subquery inside the predicate would make the engine call
    # the function one row at a time)
    assert "iterativelength(1000, __pgq_x1000.n, a.rowid + __pgq_x1000.z, b.rowid) AS len" in main
    assert "shortestpath(1000, __pgq_x1000.n, a.rowid + __pgq_x1000.z, b.rowid) AS vs" in main
    assert ("(SELECT count(__pgq_csr1000.temp) * 0 AS z, (SELECT count(*) FROM person) AS n FROM __pgq_csr1000) __pgq_x1000, "
            "person a, person b") in main
    assert "IS NOT NULL AND (a.id = 3)) gt ORDER BY dst;" in main  # MATCH keeps the pairs that have a path
    assert st[1] == "SELECT delete_csr(1000);"


def test_quantifiers_directions_and_joins():
    r = rewrite(GRAPH +
                "SELECT count(*) FROM GRAPH_TABLE (snb MATCH (a:Person)-[k:knows]->(b:Person) WHERE a.id < b.id COLUMNS (a.id, k.p2));\n"
                "SELECT * FROM GRAPH_TABLE (snb MATCH p = ANY SHORTEST (a)<-[k]-{1,3}(b) COLUMNS (path_with_edges(p) AS pe));\n"
                "SELECT * FROM GRAPH_TABLE (snb MATCH (a:Person)-[k:knows]->+(b:Person) COLUMNS (a.id, b.id));\n"
                "SELECT * FROM GRAPH_TABLE (snb MATCH (a:Person)-[k:knows]->*(b:Person) COLUMNS (a.id, b.id));\n"
                "WITH x AS (SELECT 3 AS s) SELECT * FROM x, GRAPH_TABLE (snb MATCH p = ANY SHORTEST (a)-[k]->*(b) COLUMNS (a.id AS i, "
                "cheapest_path_length(p, k.w) AS c)) g WHERE g.i = x.s;\n")
    one_hop = r[1][1]
    assert one_hop == ["SELECT count(*) FROM (SELECT a.id, k.p2 FROM person a, person b, knows k WHERE k.p1 = a.id AND k.p2 = b.id "
                       "AND (a.id < b.id)) __pgq_gt0;"]  # a join, no CSR, no cleanup
    rev = r[2][1][0]  # <-[ ]- : the path runs from b to a; {1,3} is the bounded 6-argument form
    assert "shortestpath_with_edges(1000, __pgq_x1000.n, b.rowid + __pgq_x1000.z, a.rowid) AS pe" in rev
    assert "b.rowid + __pgq_x1000.z, a.rowid, CAST(1 AS BIGINT), CAST(3 AS BIGINT)) IS NOT NULL" in rev
    plus = r[3][1][0]
    assert "CAST(1 AS BIGINT), __pgq_x1001.n) IS NOT NULL" in plus
    star = r[4][1][0]  # no path variable, no ANY SHORTEST: existence only
    assert "WHERE reachability(1002, __pgq_x1002.n, a.rowid + __pgq_x1002.z, b.rowid)) __pgq_gt0" in star
    cheap = r[5][1][0]  # joins an existing WITH list; the weight column travels with the edge rows
    assert cheap.startswith("WITH __pgq_csr1003 AS (") and "),  x AS (SELECT 3 AS s) SELECT * FROM x, (" in cheap
    assert "e.rowid, CAST(e.w AS BIGINT)) AS temp" in cheap
    assert "cheapest_path_length(1003, __pgq_x1003.n, a.rowid + __pgq_x1003.z, b.rowid) AS c" in cheap


def test_unsupported_syntax_is_reported_not_guessed():
    two = ("CREATE PROPERTY GRAPH g2 VERTEX TABLES (person, city) EDGE TABLES (knows SOURCE KEY (p1) REFERENCES person (id) "
           "DESTINATION KEY (p2) REFERENCES person (id), lives SOURCE KEY (p) REFERENCES person (id) DESTINATION KEY (c) "
           "REFERENCES city (id));\n")
    r = rewrite(GRAPH + two +
                "SELECT * FROM GRAPH_TABLE (nope MATCH (a)-[k]->(b) COLUMNS (a.id));\n"
                "SELECT * FROM GRAPH_TABLE (snb MATCH (a:Person)-[k:knows]-(b:Person) COLUMNS (a.id));\n"
                "SELECT * FROM GRAPH_TABLE (snb MATCH p = ALL SHORTEST (a)-[k]->*(b) COLUMNS (a.id));\n"
                "SELECT * FROM GRAPH_TABLE (snb MATCH p = ANY SHORTEST (a)-[k]->(b) COLUMNS (a.id));\n"
                "SELECT * FROM GRAPH_TABLE (g2 MATCH (a)-[k]->(b) COLUMNS (a.id));\n"
                "SELECT * FROM GRAPH_TABLE (g2 MATCH (a:person)-[k:lives]->*(b:city) COLUMNS (a.id));\n"
                "SELECT * FROM GRAPH_TABLE (g2 MATCH (a:city)-[k:lives]->(b:person) COLUMNS (a.id));\n"
                "SELECT * FROM GRAPH_TABLE (g2 MATCH (a:person)-[k:lives]->(c:city)<-[l:lives]-(b:person) COLUMNS (a.id, b.id, c.id));\n"
                "CREATE PROPERTY GRAPH bad VERTEX TABLES (person) EDGE TABLES (knows SOURCE KEY (p1) REFERENCES nobody (id) "
                "DESTINATION KEY (p2) REFERENCES person (id));\n")
    msgs = [x[1] if x[0] == "error" else x for x in r[2:]]
    assert "property graph nope does not exist" in msgs[0]
    assert isinstance(msgs[1], str) and msgs[1].startswith("pgq:")  # undirected edge
    assert "only ANY SHORTEST" in msgs[2]
    assert "needs a quantified edge" in msgs[3]
    assert "needs a label" in msgs[4]
    assert "same vertex table" in msgs[5]
    assert "runs from person to city" in msgs[6]
    assert msgs[7][0] == "rewritten"  # two one-hop edges over two vertex tables: plain joins
    assert "lives k, lives l WHERE k.p = a.id AND k.c = c.id AND l.p = b.id AND l.c = c.id" in msgs[7][1][0]
    assert "not a vertex table" in msgs[8]


MACROS = """CREATE MACRO create_csr_vertex(a, b, c, d) AS d;
CREATE MACRO create_csr_edge(a, b, c, d, e, f) AS 1;
CREATE MACRO iterativelength(a, b, c, d) AS (CASE WHEN c = d THEN 0 ELSE 1 END);
CREATE MACRO shortestpath(a, b, c, d) AS c * 1000 + d;
CREATE MACRO reachability(a, b, c, d) AS (c <= d);
CREATE MACRO delete_csr(a) AS TRUE;
"""


@needs_host
def test_rewritten_sql_runs_on_the_reference_engine_with_stand_in_macros():
    """No GPU here: SQL macros with the scalar functions' names stand in for them, so this checks that the rewritten
    statements parse, bind and execute on the engine built from /root/reference, and that the joins, the predicates and
    the dense-id (rowid) arguments are the intended ones."""
    sql = MACROS + """CREATE TABLE person(id BIGINT);
CREATE TABLE knows(p1 BIGINT, p2 BIGINT);
INSERT INTO person VALUES (10), (20), (30), (40);
INSERT INTO knows VALUES (10, 20), (20, 30), (30, 10), (10, 30);
""" + GRAPH + """SELECT * FROM GRAPH_TABLE (snb MATCH p = ANY SHORTEST (a:Person WHERE a.id = 20)-[k:knows]->*(b:Person)
   COLUMNS (a.id AS src, b.id AS dst, path_length(p) AS len, vertices(p) AS vs)) gt ORDER BY dst;
SELECT a_id, count(*) FROM GRAPH_TABLE (snb MATCH (a:Person)-[k:knows]->(b:Person) COLUMNS (a.id AS a_id, b.id AS b_id)) GROUP BY a_id ORDER BY a_id;
SELECT count(*) FROM GRAPH_TABLE (snb MATCH (a:Person)-[k:knows]->*(b:Person) COLUMNS (a.id, b.id));
"""
    r = run_sql(sql)
    assert [k for k, _ in r] == ["result"] * len(r), r
    got = r[-3][1]  # rowids: 10->0, 20->1, 30->2, 40->3; the stand-in "path" is src_rowid * 1000 + dst_rowid
    assert got == ["20,10,1,1000", "20,20,0,1001", "20,30,1,1002", "20,40,1,1003"]
    assert r[-2][1] == ["10,2", "20,1", "30,1"]
    assert r[-1][1] == ["10"]  # stand-in reachability: src rowid <= dst rowid, 4 + 3 + 2 + 1 pairs


@needs_host
def test_rewritten_sql_binds_against_the_real_functions():
    """The extension LOADs without a GPU (calls fail loudly, binding works), so EXPLAIN of the rewritten statements
    checks every generated call against the registered signatures: 4- and 6-argument iterativelength, reachability,
    shortestpath, shortestpath_with_edges, cheapest_path_length / _double with the 7-argument create_csr_edge."""
    qs = ["SELECT * FROM GRAPH_TABLE (snb MATCH p = ANY SHORTEST (a:Person WHERE a.id = 10)-[k:knows]->*(b:Person) "
          "COLUMNS (a.id, b.id, path_length(p) AS len, vertices(p) AS vs, path_with_edges(p) AS pe))",
          "SELECT * FROM GRAPH_TABLE (snb MATCH p = ANY SHORTEST (a:Person)-[k:knows]->{2,5}(b:Person) COLUMNS (path_length(p) AS len))",
          "SELECT * FROM GRAPH_TABLE (snb MATCH (a:Person)<-[k:knows]-+(b:Person) COLUMNS (a.id, b.id))",
          "SELECT * FROM GRAPH_TABLE (snb MATCH (a:Person)-[k:knows]->*(b:Person) COLUMNS (a.id, b.id))",
          "SELECT * FROM GRAPH_TABLE (snb MATCH p = ANY SHORTEST (a:Person)-[k:knows]->*(b:Person) COLUMNS (cheapest_path_length(p, k.w) AS c))",
          "SELECT * FROM GRAPH_TABLE (snb MATCH p = ANY SHORTEST (a:Person)-[k:knows]->*(b:Person) "
          "COLUMNS (cheapest_path_length_double(p, k.wd) AS c))"]
    sql = ("LOAD '%s';\nCREATE TABLE person(id BIGINT);\nCREATE TABLE knows(p1 BIGINT, p2 BIGINT, w BIGINT, wd DOUBLE);\n" % EXT
           + GRAPH + "".join("EXPLAIN " + q + ";\n" for q in qs))
    r = run_sql(sql)
    assert [k for k, _ in r] == ["result"] * len(r), [x for x in r if x[0] != "result"]
    assert all(len(rows) > 5 for _, rows in r[4:])  # a plan came back for every statement


@needs_host
def test_pgq_rewrite_scalar_function_of_the_extension():
    """Clients that cannot hook the parser ask the extension: SELECT pgq_rewrite(statement) returns what to run."""
    q = "SELECT * FROM GRAPH_TABLE (g MATCH p = ANY SHORTEST (a)-[k]->*(b) COLUMNS (path_length(p) AS len))"
    sql = ("LOAD '%s';\n" % EXT +
           "SELECT pgq_rewrite('CREATE PROPERTY GRAPH g VERTEX TABLES (person) EDGE TABLES (knows SOURCE KEY (p1) REFERENCES "
           "person (id) DESTINATION KEY (p2) REFERENCES person (id))');\n"
           "SELECT pgq_rewrite('%s');\nSELECT pgq_rewrite('SELECT 42');\nSELECT pgq_rewrite(NULL);\n"
           "SELECT pgq_rewrite('SELECT * FROM GRAPH_TABLE (nope MATCH (a)-[k]->(b) COLUMNS (a.id))');\n" % q)
    r = run_sql(sql)
    assert r[1] == ("result", [""])
    assert r[2][0] == "result" and len(r[2][1]) == 2
    assert r[2][1][0].startswith("WITH __pgq_csr") and "iterativelength(" in r[2][1][0]
    assert r[2][1][1].startswith("SELECT delete_csr(")
    assert "person a, person b WHERE iterativelength(" in r[2][1][0]
    assert r[3] == ("result", ["SELECT 42"])
    assert r[4] == ("result", ["NULL"])
    assert r[5][0] == "error" and "does not exist" in r[5][1]


@pytest.mark.gpu
@needs_host
def test_graph_table_queries_match_the_oracle(tmp_path):
    """The real thing: CREATE PROPERTY GRAPH + GRAPH_TABLE queries through pgq_sql (rewriter -> reference engine ->
    extension -> GPU), answers against the oracle on dense ids."""
    V, E = 600, 5000
    src, dst = orc.gen_knows(1234, V, E)
    v, e, eids = orc.csr_build(V, src, dst)
    w = np.random.default_rng(3).integers(1, 50, size=E).astype(np.int64)
    wc = orc.csr_weights(V, v, src, w)
    ids = np.arange(V) * 7 + 3  # sparse ids: dense id = rowid
    _write_csv(tmp_path / "person.csv", [ids])
    _write_csv(tmp_path / "knows.csv", [ids[src], ids[dst], w])
    nsrc = 40
    sql = """LOAD '%s';
PRAGMA threads=1;
CREATE TABLE person(id BIGINT);
CREATE TABLE knows(p1 BIGINT, p2 BIGINT, w BIGINT);
COPY person FROM '%s';
COPY knows FROM '%s';
CREATE PROPERTY GRAPH snb VERTEX TABLES (person KEY (id) LABEL Person)
  EDGE TABLES (knows SOURCE KEY (p1) REFERENCES person (id) DESTINATION KEY (p2) REFERENCES person (id) LABEL knows);
SELECT src, dst, len, vs, pe, c FROM GRAPH_TABLE (snb MATCH p = ANY SHORTEST (a:Person WHERE a.id < %d)-[k:knows]->*(b:Person)
   COLUMNS (a.id AS src, b.id AS dst, path_length(p) AS len, vertices(p) AS vs, path_with_edges(p) AS pe,
            cheapest_path_length(p, k.w) AS c)) gt ORDER BY src, dst;
SELECT src, dst, len FROM GRAPH_TABLE (snb MATCH p = ANY SHORTEST (a:Person WHERE a.id < %d)-[k:knows]->{2,3}(b:Person)
   COLUMNS (a.id AS src, b.id AS dst, path_length(p) AS len)) ORDER BY src, dst;
SELECT count(*) FROM GRAPH_TABLE (snb MATCH (a:Person WHERE a.id < %d)-[k:knows]->*(b:Person) COLUMNS (a.id, b.id));
SELECT count(*), sum(k_w) FROM GRAPH_TABLE (snb MATCH (a:Person)-[k:knows]->(b:Person) COLUMNS (k.w AS k_w));
SELECT src, dst, len FROM GRAPH_TABLE (snb MATCH p = ANY SHORTEST (a:Person)<-[k:knows]-+(b:Person WHERE b.id < %d)
   COLUMNS (a.id AS src, b.id AS dst, path_length(p) AS len)) ORDER BY dst, src;
SELECT iterativelength(1000, CAST(%d AS BIGINT), CAST(1 AS BIGINT), CAST(2 AS BIGINT));
""" % (EXT, tmp_path / "person.csv", tmp_path / "knows.csv", nsrc * 7 + 3, nsrc * 7 + 3, nsrc * 7 + 3, nsrc * 7 + 3, V)
    r = run_sql(sql)
    assert [k for k, _ in r[:7]] == ["result"] * 7, r[:8]
    ps = np.repeat(np.arange(nsrc), V).astype(np.int64)
    pd = np.tile(np.arange(V), nsrc).astype(np.int64)
    lo, lv = orc.iterativelength(V, v, e, ps, pd)
    off, ln, pv, vals = orc.shortestpath(V, v, e, eids, ps, pd, mode=0)
    offe, lne, pve, valse = orc.shortestpath(V, v, e, eids, ps, pd, mode=1)
    co, cv, _ = orc.cheapest_path_length(V, v, e, wc, ps, pd)
    # 1: every reachable (src, dst) pair, its hop count, one shortest path (vertex rowids; with edge rowids) and its cost
    assert r[7][0] == "result", r[7]
    exp = []
    for i in range(len(ps)):
        if not lv[i]:
            continue
        vs = [int(x) for x in vals[int(off[i]):int(off[i]) + int(ln[i])]]
        pe = [int(x) for x in valse[int(offe[i]):int(offe[i]) + int(lne[i])]]
        exp.append((int(ids[ps[i]]), int(ids[pd[i]]), int(lo[i]), vs, pe, int(co[i])))
    got = []
    for row in r[7][1]:
        a, b, ln_, rest = row.split(",", 3)
        vs_s, rest = rest.split("],", 1)
        pe_s, c_s = rest.rsplit(",", 1)
        got.append((int(a), int(b), int(ln_), _parse_list(vs_s + "]"), _parse_list(pe_s), int(c_s)))
    assert len(got) == len(exp) and got == exp
    # 2: {2,3}: only pairs whose shortest path has 2 or 3 hops
    bo, bv = orc.iterativelength_bounded(V, v, e, ps, pd, 2, 3)
    exp = ["%d,%d,%d" % (ids[ps[i]], ids[pd[i]], bo[i]) for i in range(len(ps)) if bv[i]]
    assert r[8] == ("result", exp)
    # 3: existence only
    assert r[9] == ("result", [str(int(lv.sum()))])
    # 4: one hop = the edge table itself
    assert r[10] == ("result", ["%d,%d" % (E, int(w.sum()))])
    # 5: reversed edges, at least one hop: paths run from b to a
    bo, bv = orc.iterativelength_bounded(V, v, e, ps, pd, 1, V)
    exp = ["%d,%d,%d" % (ids[pd[i]], ids[ps[i]], bo[i]) for i in range(len(ps)) if bv[i]]
    assert r[11] == ("result", exp)
    # the CSRs the statements built are gone again
    assert r[12][0] == "error" and "does not exist" in r[12][1]
