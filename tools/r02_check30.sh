#!/bin/bash
# round-2 GPU check 30: cheapest_path_length tests (incl. hub chunks on 32-bit rows) and the default bench line
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -q -x -p no:faulthandler -k "cheapest" 2>&1 | tail -3
timeout 1700 python bench.py > gpurun_out/c30_bench.json 2> gpurun_out/c30_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/c30_bench.json'))
p=d['paths']
r=d['roofline']
print("BENCH", round(d['value']), round(d['ms_per_step'],1), 'e2e', round(d['e2e']['value']), 'frac', round(r['frac'],3), 'traffic', r['traffic'], 'dram_frac', r.get('dram_frac'),
  'csr', round(d['csr_build']['value'],1), d['csr_build']['roofline'].get('dram_frac'),
  'C4', round(p['shortestpath']['pairs_per_s']), round(p['iterativelength']['pairs_per_s']), round(p['cheapest_path_length']['pairs_per_s']), round(p['cheapest_path_length']['rows_32_bit']['pairs_per_s']))
print(d['parity']); print(d['sql'])
PY
