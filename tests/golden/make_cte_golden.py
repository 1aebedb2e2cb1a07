"""Generates tests/golden/cte_lengths.json: hop distances computed by the REFERENCE DuckDB
(built from /root/reference, shell at baseline/_ref/duckdb/duckdb) with the recursive-CTE BFS
of /root/reference/benchmark/ldbc/queries/bi-10-shortestpath.sql:8-31, on a seeded R-MAT graph.
The oracle's iterativelength must reproduce them (tests/test_oracle_cpu.py).  Run in the build
container only (the GPU box has no /root/reference and needs only the committed JSON)."""
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
SCALE, SEED, NSRC, NDST = 9, 0x5eed0000 + 9, 24, 20

src, dst = orc.gen_rmat(SEED, SCALE, E=6 << SCALE)
V = 1 << SCALE
v, e, _ = orc.csr_build(V, src, dst)
qs_all, qd_all = orc.gen_pairs(42, V, v, NSRC * NDST)
sources = sorted(set(int(x) for x in qs_all[:NSRC]))

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
     WHERE f.friend = k.k_person1id AND f.hopCount < 16
), friends_shortest AS (
    SELECT startPerson, min(hopCount) AS hopCount, friend FROM friends WHERE hopCount > 0
     GROUP BY startPerson, friend
)
SELECT startPerson, friend, hopCount FROM friends_shortest ORDER BY 1,2;
""" % (csv, ",".join("(%d)" % s for s in sources))
    out = subprocess.run([DUCKDB], input=sql, capture_output=True, text=True, check=True).stdout

dist = {}
for line in out.strip().splitlines():
    parts = line.strip().split(",")
    if len(parts) == 3 and parts[0].isdigit():
        dist[(int(parts[0]), int(parts[1]))] = int(parts[2])

q_src, q_dst, expected = [], [], []
for i, s in enumerate(sources):
    for j in range(NDST):
        d = int(qd_all[(i * NDST + j) % len(qd_all)])
        q_src.append(s)
        q_dst.append(d)
        if s == d:
            expected.append(0)   # iterativelength: src==dst -> 0 without a search
        else:
            expected.append(dist.get((s, d)))  # None = unreachable = SQL NULL
json.dump(dict(note="hop distances from the reference DuckDB recursive CTE; see make_cte_golden.py",
               V=V, src=[int(x) for x in src], dst=[int(x) for x in dst],
               q_src=q_src, q_dst=q_dst, expected=expected),
          open(os.path.join(ROOT, "tests", "golden", "cte_lengths.json"), "w"))
print("pairs", len(expected), "reachable", sum(x is not None for x in expected),
      "hist", np.bincount([x for x in expected if x is not None]).tolist())
