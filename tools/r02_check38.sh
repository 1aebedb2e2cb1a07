#!/bin/bash
# round-2 GPU check 38: a pairs table large enough for a parallel scan (1 M pairs): PRAGMA threads 1 against 8, batcher stats
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for t in 1 8; do
  timeout 1200 python tools/sql_bench.py --persons 100000 --knows 4000000 --pairs 1000000 --threads $t --reps 2 > gpurun_out/c38_t$t.json 2> gpurun_out/c38_t$t.err
  echo "threads $t rc=$?"; tail -2 gpurun_out/c38_t$t.err | cut -c1-300
  python - <<PY
import json
d=json.load(open('gpurun_out/c38_t$t.json'))
print({k:round(v,1) for k,v in d['statements_ms'].items() if v is not None and k.endswith('_1') and 'subquery' not in k}, d['got'].get('batcher_0'), d['got'].get('batcher_1'), d['parity_vs_c_abi'], d['errors'])
PY
done
