#!/bin/bash
# round-2 GPU check 21: the SQL/PGQ rewriter end to end + C4 per-call samples
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_pgq_rewrite.py tests/test_sql_extension.py -x -q -p no:faulthandler > gpurun_out/c21_pytest.log 2>&1
echo "pytest rc=$?"; tail -30 gpurun_out/c21_pytest.log | cut -c1-400
timeout 900 python bench.py --steps 3 --warmup 3 --no-csr > gpurun_out/c21_bench.json 2> gpurun_out/c21_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/c21_bench.json'))
p=d['paths']
for k in ('shortestpath','iterativelength'):
    print(k, round(p[k]['pairs_per_s']), p[k]['ms_samples'])
print('cheap', {k:v for k,v in p['cheapest_path_length'].items() if k in ('pairs_per_s','ms','rounds','parity')})
print(p['cheapest_path_length']['roofline'])
PY
