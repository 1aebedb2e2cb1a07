#!/bin/bash
# round-2 GPU check 37: the SQL path at C4 size (450 k persons, 40 M knows rows, 16 384 pairs), PRAGMA threads 1 and 8
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for t in 1 8; do
  T0=$(date +%s)
  timeout 1500 python tools/sql_bench.py --persons 450000 --knows 40000000 --pairs 16384 --threads $t --reps 3 > gpurun_out/c37_sql_c4_t$t.json 2> gpurun_out/c37_sql_c4_t$t.err
  echo "threads $t rc=$? wall=$(( $(date +%s) - T0 )) s"; tail -2 gpurun_out/c37_sql_c4_t$t.err | cut -c1-300
  python - <<PY
import json
d=json.load(open('gpurun_out/c37_sql_c4_t$t.json'))
print({k:round(v,1) for k,v in d['statements_ms'].items() if v is not None and (k.endswith('_2') or k=='copy2')}, d['graph_table'], d['parity_vs_c_abi'], d['errors'])
PY
done
