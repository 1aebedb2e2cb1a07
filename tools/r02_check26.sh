#!/bin/bash
# round-2 GPU check 26: the full GPU suite, smoke and the default bench line on the final tree
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q -p no:faulthandler > gpurun_out/c26_pytest.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/c26_pytest.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c26_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/c26_smoke.log
timeout 1700 python bench.py > gpurun_out/c26_bench.json 2> gpurun_out/c26_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/c26_bench.json'))
p=d['paths']
print("BENCH", round(d['value']), round(d['ms_per_step'],1), 'e2e', round(d['e2e']['value']), 'frac', round(d['roofline']['frac'],3), 'traffic', d['roofline']['traffic'] is not None,
  'csr', round(d['csr_build']['value'],1), d['csr_build']['roofline'].get('traffic') is not None,
  'C4', round(p['shortestpath']['pairs_per_s']), round(p['iterativelength']['pairs_per_s']), round(p['cheapest_path_length']['pairs_per_s']))
print(d['parity'])
PY
