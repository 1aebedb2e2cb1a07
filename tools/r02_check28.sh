#!/bin/bash
# round-2 GPU check 28: CUB SortPairs as a reference point for the CSR build's sort; bench sql section
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python tools/cubref/run.py 2>&1 | tail -4
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-csr > gpurun_out/c28_bench.json 2> gpurun_out/c28_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/c28_bench.json'))
print(json.dumps(d['sql'])[:1500])
PY
