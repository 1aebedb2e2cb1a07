#!/bin/bash
# round-2 GPU check 16: cheapest_path_length on 16-bit rows with a destination bound, batches in flight
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -q -x -p no:faulthandler -k "cheapest" > gpurun_out/c16_pytest.log 2>&1
echo "pytest rc=$?"; tail -15 gpurun_out/c16_pytest.log | cut -c1-300
timeout 900 python -m pytest tests/test_sql_extension.py tests/test_pgq_rewrite.py -q -x -p no:faulthandler > gpurun_out/c16_pytest_sql.log 2>&1
echo "pytest sql rc=$?"; tail -5 gpurun_out/c16_pytest_sql.log | cut -c1-300
timeout 900 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-csr --pairs 1024 > gpurun_out/c16_bench.json 2> gpurun_out/c16_bench.err
echo "bench rc=$?"; tail -3 gpurun_out/c16_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/c16_bench.json'))
print(json.dumps(d['paths']['cheapest_path_length'], indent=1))
PY
