#!/bin/bash
# round-2 GPU check 27: the bench line's sql section
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-csr > gpurun_out/c27_bench.json 2> gpurun_out/c27_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/c27_bench.json'))
print(json.dumps(d['sql'], indent=1))
PY
