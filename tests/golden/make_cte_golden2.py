"""Generates tests/golden/cte_knows_bounded.json: hop distances with an UPPER BOUND on the path length, computed
by the REFERENCE DuckDB (built from /root/reference, shell at baseline/_ref/duckdb/duckdb) with the recursive-CTE
BFS of /root/reference/benchmark/ldbc/queries/bi-10-shortestpath.sql:8-31 (its `hopCount < N` guard is the bound),
on a seeded LDBC-knows-shaped graph (both directions materialised, as /root/reference/benchmark/ldbc/
snb-load.sql:23-24 does).  Pins iterativelength, reachability and the upper bound of the 6-argument iterativelength
(rule R11; the LOWER bound cannot be pinned this way: the CTE enumerates walks, the function reports the BFS level).
Run in the build container only; the GPU box needs only the committed JSON."""
import json
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import orc

DUCKDB = os.path.join(ROOT, "baseline", "_ref", "duckdb", "duckdb")
V, E, SEED, NSRC, NDST = 700, 4200, 2024, 16, 30
UPPERS = [1, 2, 3, 4, 24]

src, dst = orc.gen_knows(SEED, V, E)
v, e, _ = orc.csr_build(V, src, dst)
qs_all, qd_all = orc.gen_pairs(7, V, v, NSRC * NDST)
sources = sorted(set(int(x) for x in qs_all[:NSRC]))


def run_cte(upper):
    with tempfile.TemporaryDirectory() as td:
        csv = os.path.join(td, "knows.csv")
        with open(csv, "w") as f:
            for a, b in zip(src, dst):
                f.write("%d,%d\n" % (a, b))
        sql = """
CREATE TABLE knows(k_person1id BIGINT, k_person2id BIGINT);
COPY knows FROM '%s';
CREATE TABLE starts(p BIGINT);
INSERT INTO starts VALUES %s;
.mode csv
WITH RECURSIVE friends(startPerson, hopCount, friend) AS (
    SELECT p, 0, p FROM starts
  UNION
    SELECT f.startPerson, f.hopCount+1, k.k_person2id
      FROM friends f, knows k
     WHERE f.friend = k.k_person1id AND f.hopCount < %d
), friends_shortest AS (
    SELECT startPerson, min(hopCount) AS hopCount, friend FROM friends WHERE hopCount > 0
     GROUP BY startPerson, friend
)
SELECT startPerson, friend, hopCount FROM friends_shortest ORDER BY 1,2;
""" % (csv, ",".join("(%d)" % s for s in sources), upper)
        out = subprocess.run([DUCKDB], input=sql, capture_output=True, text=True, check=True).stdout
    dist = {}
    for line in out.strip().splitlines():
        parts = line.strip().split(",")
        if len(parts) == 3 and parts[0].isdigit():
            dist[(int(parts[0]), int(parts[1]))] = int(parts[2])
    return dist


q_src, q_dst = [], []
for i, s in enumerate(sources):
    for j in range(NDST):
        q_src.append(s)
        q_dst.append(int(qd_all[(i * NDST + j) % len(qd_all)]))
expected = {}
for u in UPPERS:
    dist = run_cte(u)
    expected[str(u)] = [0 if s == d else dist.get((s, d)) for s, d in zip(q_src, q_dst)]  # None = SQL NULL
    print("upper", u, "reachable", sum(x is not None for x in expected[str(u)]),
          "hist", np.bincount([x for x in expected[str(u)] if x is not None]).tolist())
json.dump(dict(note="hop distances with an upper bound, from the reference DuckDB recursive CTE; see make_cte_golden2.py",
               V=V, src=[int(x) for x in src], dst=[int(x) for x in dst], q_src=q_src, q_dst=q_dst,
               uppers=UPPERS, expected=expected),
          open(os.path.join(ROOT, "tests", "golden", "cte_knows_bounded.json"), "w"))
