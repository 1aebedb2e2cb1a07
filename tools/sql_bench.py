"""SQL-path timing: the PGQ functions called by the reference DuckDB through the loadable extension.

  python tools/sql_bench.py [--persons 10000 --knows 180000 --pairs 2048] [--threads 1]

Generates an LDBC-knows-shaped person/knows/pairs table set (device generator, same as bench.py),
loads it into the reference DuckDB (baseline/_ref/duckdb/libduckdb.so) with the pgq_sql driver,
LOADs pgq_b200.duckdb_extension and times each statement (Connection::Query wall time: parse, plan,
the engine's joins/aggregates, our functions, result materialisation).  Prints one JSON line.
Needs a GPU.  Not part of the product."""
import argparse
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pgq_b200  # noqa: E402
from bench import SEED_BASE, knows_params, make_pairs  # noqa: E402

EXT = os.path.join(ROOT, "duckdb-pgq.old_b200", "ext", "pgq_b200.duckdb_extension")
DRIVER = os.path.join(ROOT, "duckdb-pgq.old_b200", "ext", "pgq_sql")


def write_csv(path, cols):
    if len(cols[0]) > 1000000:  # np.savetxt takes minutes for tens of millions of rows
        import pandas as pd
        pd.DataFrame({str(i): c for i, c in enumerate(cols)}).to_csv(path, header=False, index=False)
        return
    np.savetxt(path, np.stack(cols, axis=1), fmt="%d", delimiter=",")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--persons", type=int, default=10000)
    ap.add_argument("--knows", type=int, default=180000)
    ap.add_argument("--pairs", type=int, default=2048)
    ap.add_argument("--threads", type=int, default=1)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    V, E, n = a.persons, a.knows, a.pairs
    ctx = pgq_b200.Context(0)
    mult, add = knows_params(V)
    d_src, d_dst = ctx.dev_alloc(E * 8), ctx.dev_alloc(E * 8)
    ctx.gen_knows_device(SEED_BASE + 1, V, E, mult, add, d_src, d_dst)
    src, dst = np.empty(E, dtype=np.int64), np.empty(E, dtype=np.int64)
    ctx.d2h(src, d_src)
    ctx.d2h(dst, d_dst)
    ctx.dev_free(d_src)
    ctx.dev_free(d_dst)
    deg = np.bincount(src, minlength=V).astype(np.int64)
    v_off = np.concatenate([[0], np.cumsum(deg), [E]]).astype(np.int64)
    ps, pd = make_pairs(V, v_off, n, 77)
    # the C-ABI answer for the same pairs (checksum the SQL answer must reproduce)
    ctx.csr_build(900, V, src, dst)
    lo, lv = ctx.iterativelength(900, ps, pd)
    want_sum, want_cnt = int(lo[lv == 1].sum()), int((lv == 1).sum())
    # SQL/PGQ form (pgq_rewrite.hpp): 16 sources x every person
    nsrc = 16
    gs = np.repeat(np.arange(nsrc, dtype=np.int64), V)
    gd = np.tile(np.arange(V, dtype=np.int64), nsrc)
    glo, glv = ctx.iterativelength(900, gs, gd)
    gt_sum, gt_cnt = int(glo[glv == 1].sum()), int((glv == 1).sum())
    ctx.csr_delete(900)
    ctx.close()
    tmp = tempfile.mkdtemp(prefix="pgq_sql_")
    write_csv(os.path.join(tmp, "person.csv"), [np.arange(V, dtype=np.int64) * 7 + 3])
    write_csv(os.path.join(tmp, "knows.csv"), [src * 7 + 3, dst * 7 + 3])
    write_csv(os.path.join(tmp, "pairs.csv"), [np.arange(n, dtype=np.int64), ps, pd])
    vq = "(SELECT count(*) FROM person)"
    build = """SELECT sum(create_csr_edge(0, %s,
    CAST((SELECT sum(create_csr_vertex(0, %s, sub.dense_id, sub.cnt))
          FROM (SELECT p.rowid AS dense_id, count(k.p1) AS cnt FROM person p LEFT JOIN knows k ON k.p1 = p.id
                GROUP BY p.rowid) sub) AS BIGINT),
    a.rowid, b.rowid, k.rowid))
  FROM knows k JOIN person a ON k.p1 = a.id JOIN person b ON k.p2 = b.id;""" % (vq, vq)
    stmts = [("load", "LOAD '%s';" % EXT), ("threads", "PRAGMA threads=%d;" % a.threads),
             ("ddl1", "CREATE TABLE person(id BIGINT);"), ("ddl2", "CREATE TABLE knows(p1 BIGINT, p2 BIGINT);"),
             ("ddl3", "CREATE TABLE pairs(i BIGINT, s BIGINT, d BIGINT);"),
             ("copy1", "COPY person FROM '%s';" % os.path.join(tmp, "person.csv")),
             ("copy2", "COPY knows FROM '%s';" % os.path.join(tmp, "knows.csv")),
             ("copy3", "COPY pairs FROM '%s';" % os.path.join(tmp, "pairs.csv"))]
    stmts.append(("pg", "CREATE PROPERTY GRAPH snb VERTEX TABLES (person KEY (id) LABEL Person) EDGE TABLES (knows SOURCE KEY (p1) "
                        "REFERENCES person (id) DESTINATION KEY (p2) REFERENCES person (id) LABEL knows);"))
    for r in range(a.reps):
        # CREATE PROPERTY GRAPH + GRAPH_TABLE through the rewriter: CSR build (CTE), 16 x V path lengths, delete_csr
        stmts.append(("graph_table_%d" % r,
                      "SELECT sum(len), count(len) FROM GRAPH_TABLE (snb MATCH p = ANY SHORTEST (a:Person WHERE a.rowid < %d)"
                      "-[k:knows]->*(b:Person) COLUMNS (path_length(p) AS len)) g;" % nsrc))
        stmts += [("csr_build_%d" % r, build),
                  ("iterativelength_%d" % r, "SELECT sum(iterativelength(0, %s, s, d)), count(iterativelength(0, %s, s, d)) FROM pairs;" % (vq, vq)),
                  ("iterativelength_const_%d" % r,
                   "SELECT sum(iterativelength(0, CAST(%d AS BIGINT), s, d)) FROM pairs;" % V),
                  ("batcher_%d" % r, "SELECT pgq_batcher_stats(0);"),
                  ("reachability_%d" % r, "SELECT sum(CASE WHEN reachability(0, %s, s, d) THEN 1 ELSE 0 END) FROM pairs;" % vq),
                  # predicate with a constant v_size: a vectorised filter (1024-row calls)
                  ("reachability_where_const_%d" % r,
                   "SELECT count(*) FROM pairs WHERE reachability(0, CAST(%d AS BIGINT), s, d);" % V),
                  # predicate with the scalar subquery inside: the engine plans a BLOCKWISE_NL_JOIN and
                  # evaluates the predicate one row per call
                  ("reachability_where_subquery_%d" % r, "SELECT count(*) FROM pairs WHERE reachability(0, %s, s, d);" % vq)
                  if n <= 50000 and not os.environ.get("SQL_BENCH_SKIP_NL") else ("skipped_%d" % r, "SELECT 1;"),  # one row per call: minutes for a large pairs table
                  ("shortestpath_%d" % r, "SELECT count(shortestpath(0, %s, s, d)) FROM pairs;" % vq),
                  ("delete_%d" % r, "SELECT delete_csr(0);")]
    env = dict(os.environ, PGQ_SQL_TIME="1")
    p = subprocess.run([DRIVER], input="\n".join(s for _, s in stmts) + "\n", capture_output=True, text=True, env=env,
                       timeout=3000)
    import shutil
    shutil.rmtree(tmp, ignore_errors=True)
    times, results, errors, cur = {}, {}, {}, None
    for ln in p.stdout.splitlines():
        if ln.startswith("-- time "):
            k, ms = ln[8:].split(":")
            times[int(k)] = float(ms.split()[0])
        elif ln.startswith("-- result "):
            cur = int(ln[10:])
            results[cur] = []
        elif ln.startswith("-- error "):
            k, msg = ln[9:].split(":", 1)
            errors[int(k)] = msg.strip()
            cur = None
        elif ln.startswith("-- done"):
            cur = None
        elif cur is not None:
            results[cur].append(ln)
    out = {"persons": V, "knows": E, "pairs": n, "threads": a.threads, "statements_ms": {}, "errors": errors,
           "expect": {"sum": want_sum, "count": want_cnt}, "got": {}}
    for i, (name, _) in enumerate(stmts, start=1):
        if name.startswith(("csr_build", "iterativelength", "reachability", "shortestpath", "copy2", "graph_table")):
            out["statements_ms"][name] = times.get(i)
        if name.startswith(("iterativelength", "reachability", "shortestpath", "csr_build", "graph_table", "batcher")):
            out["got"][name] = results.get(i)
    last = a.reps - 1
    out["graph_table"] = {"pairs": nsrc * V, "expect": "%d,%d" % (gt_sum, gt_cnt),
                          "parity_vs_c_abi": out["got"].get("graph_table_%d" % last) == ["%d,%d" % (gt_sum, gt_cnt)]}
    out["parity_vs_c_abi"] = bool(results and out["got"].get("iterativelength_%d" % last) == ["%d,%d" % (want_sum, want_cnt)])
    print(json.dumps(out))
    if p.returncode != 0 or errors:
        sys.stderr.write(p.stderr[-2000:])
        sys.exit(1)


if __name__ == "__main__":
    main()
