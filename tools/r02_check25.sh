#!/bin/bash
# round-2 GPU check 25: batches in flight (streams) sweep for the C3 step
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for s in 8 12 16 6 8; do
  timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-csr --no-paths --streams $s > gpurun_out/c25_s$s.json 2> gpurun_out/c25_s$s.err
  python - <<PY
import json
d=json.load(open('gpurun_out/c25_s$s.json'))
print("STREAMS $s", round(d['value']), round(d['ms_per_step'],1), 'e2e', round(d['e2e']['value']), 'frac', round(d['roofline']['frac'],3))
PY
done
