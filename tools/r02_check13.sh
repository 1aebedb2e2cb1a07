#!/bin/bash
# round-2 GPU check 13: 12-bit wide sort (23/24-bit keys): parity + s24 build time; full bench rerun (paths anomaly)
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -p no:faulthandler -k "general and (csr_both_sort_cores or csr_bit_identical or large_scale)" > gpurun_out/c13_pytest.log 2>&1
echo "pytest rc=$?" | tee -a gpurun_out/c13_pytest.log
tail -8 gpurun_out/c13_pytest.log | cut -c1-250
timeout 300 python tools/csr_only.py 24 3 > gpurun_out/c13_csr24.log 2>&1; cat gpurun_out/c13_csr24.log
timeout 300 python tools/csr_only.py 24 2 sort_wide=0 > gpurun_out/c13_csr24_8bit.log 2>&1; cat gpurun_out/c13_csr24_8bit.log
timeout 300 python tools/csr_only.py 23 2 > gpurun_out/c13_csr23.log 2>&1; cat gpurun_out/c13_csr23.log
timeout 1500 python bench.py --steps 5 --warmup 3 > gpurun_out/c13_bench.json 2> gpurun_out/c13_bench.err
echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/c13_bench.json'))
p=d['paths']
print("BENCH", round(d['value']), round(d['ms_per_step'],1), 'e2e', round(d['e2e']['value']), 'chunked', round(d['e2e_chunked']['value']), 'frac', round(d['roofline']['frac'],3),
      'C4 sp', round(p['shortestpath']['pairs_per_s']), round(p['shortestpath']['e2e_chunked']['pairs_per_s']), 'len', round(p['iterativelength']['pairs_per_s']), 'c1', round(d['config1']['gpu']['pairs_per_s']), d['setup'], d['parity'])
PY
