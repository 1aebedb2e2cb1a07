#!/bin/bash
# round-2 GPU check 19: full GPU suite + smoke + quick bench after the control-block line split
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:faulthandler > gpurun_out/c19_pytest.log 2>&1
echo "pytest rc=$?"; tail -5 gpurun_out/c19_pytest.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c19_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/c19_smoke.log
for i in 1 2; do
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/c19_bench_$i.json 2> gpurun_out/c19_bench_$i.err
echo "bench rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/c19_bench_$i.json'))
p=d['paths']; c=d['config1']
print("BENCH", round(d['value']), round(d['ms_per_step'],1), 'e2e', round(d['e2e']['value']), 'chunked', round(d['e2e_chunked']['value']), 'frac', round(d['roofline']['frac'],3),
  'C4 sp', round(p['shortestpath']['pairs_per_s']), 'len', round(p['iterativelength']['pairs_per_s']), 'cheap', round(p['cheapest_path_length']['pairs_per_s']),
  'c1', round(c['value']) if 'value' in c else c.keys(), 'csr', round(d['csr_build']['value'],1))
print(d['parity'])
PY
done
