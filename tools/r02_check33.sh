#!/bin/bash
# round-2 GPU check 33: cudaLimitMaxL2FetchGranularity 32 / 64 / 128 for the C3 step and the C4 calls
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for g in default 32 64 128 default; do
  if [ $g = default ]; then unset PGQ_B200_L2_FETCH; else export PGQ_B200_L2_FETCH=$g; fi
  PGQ_B200_VERBOSE=1 timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-csr > gpurun_out/c33_$g.json 2> gpurun_out/c33_$g.err
  grep -m1 "L2 fetch" gpurun_out/c33_$g.err
  python - <<PY
import json
d=json.load(open('gpurun_out/c33_$g.json'))
p=d['paths']
print("FETCH $g", round(d['value']), round(d['ms_per_step'],1), 'C4', round(p['shortestpath']['pairs_per_s']), round(p['iterativelength']['pairs_per_s']), round(p['cheapest_path_length']['pairs_per_s']), round(p['cheapest_path_length']['double_weights_fractional']['pairs_per_s']))
PY
done
