#!/bin/bash
# round-2 GPU check 40: pattern of the slow shortestpath statements on the C1-shaped SQL run (with / without the 2048 one-row calls before them)
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for skip in "" 1 "" 1; do
SQL_BENCH_SKIP_NL=$skip timeout 600 python tools/sql_bench.py --reps 8 > gpurun_out/c40.json 2>/dev/null
python - <<PY
import json
d=json.load(open('gpurun_out/c40.json'))
print("skip_nl='$skip' shortestpath", [round(d['statements_ms']['shortestpath_%d'%k],2) for k in range(8)])
PY
done
