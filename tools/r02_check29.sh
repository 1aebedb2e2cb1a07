#!/bin/bash
# round-2 GPU check 29: is the 16 ms shortestpath statement of check 28 reproducible?
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for i in 1 2 3 4; do
  timeout 600 python tools/sql_bench.py --reps 4 > gpurun_out/c29_$i.json 2>/dev/null
  python - <<PY
import json
d=json.load(open('gpurun_out/c29_$i.json'))
print({k:round(v,2) for k,v in d['statements_ms'].items() if k.startswith(('shortestpath','iterativelength','graph_table'))})
PY
done
