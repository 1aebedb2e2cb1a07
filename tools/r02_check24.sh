#!/bin/bash
# round-2 GPU check 24: SQL/PGQ rewriter end to end after the |V|-as-a-column change; SQL-path timings incl. GRAPH_TABLE
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_pgq_rewrite.py tests/test_sql_extension.py -x -q -p no:faulthandler > gpurun_out/c24_pytest.log 2>&1
echo "pytest rc=$?"; tail -5 gpurun_out/c24_pytest.log | cut -c1-400
timeout 900 python tools/sql_bench.py > gpurun_out/c24_sql_c1.json 2> gpurun_out/c24_sql_c1.err; echo "sql c1 rc=$?"; tail -3 gpurun_out/c24_sql_c1.err
timeout 1200 python tools/sql_bench.py --persons 100000 --knows 4000000 --pairs 16384 > gpurun_out/c24_sql_100k.json 2> gpurun_out/c24_sql_100k.err; echo "sql 100k rc=$?"; tail -3 gpurun_out/c24_sql_100k.err
timeout 900 python tools/sql_bench.py --threads 4 > gpurun_out/c24_sql_c1_t4.json 2> gpurun_out/c24_sql_c1_t4.err; echo "sql c1 t4 rc=$?"
python - <<'PY'
import json
for f in ('c24_sql_c1','c24_sql_100k','c24_sql_c1_t4'):
    try:
        d=json.load(open('gpurun_out/%s.json'%f))
        print(f, {k:round(v,2) for k,v in d['statements_ms'].items() if k.endswith('_2') or k=='copy2'}, d['graph_table'], d['parity_vs_c_abi'], d['errors'])
    except Exception as ex: print(f, 'ERR', ex)
PY
