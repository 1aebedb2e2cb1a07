#!/bin/bash
# round-2 GPU check 11: C4 paths slowdown seen in check 10; dense default vs sparse+done; rest of the suite
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python tools/paths_trace.py 2>&1 | grep -v "^\[pgq\] iter" > gpurun_out/c11_paths_trace.log
cat gpurun_out/c11_paths_trace.log
for v in "" "dense_store=0 done_bitmap=1"; do
  echo "== BASE $v" >> gpurun_out/c11_ab.log
  timeout 300 python tools/trace_levels.py 24 512 $v 2>&1 | grep "iter [1-8]\|level_kernel_ms\|rror" | sed -E 's/next:.*//; s/.*level_kernel_ms.: ([0-9.]+).*/total \1/' >> gpurun_out/c11_ab.log
done
cat gpurun_out/c11_ab.log
timeout 1800 python -m pytest tests/test_bench_contract.py tests/test_sql_extension.py tests/test_zz_no_alloc.py -m gpu -x -q -p no:faulthandler > gpurun_out/c11_pytest.log 2>&1
echo "pytest rc=$?" | tee -a gpurun_out/c11_pytest.log
tail -6 gpurun_out/c11_pytest.log | cut -c1-250
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c11_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/c11_smoke.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/c11_smoke_launches.csv python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c11_smoke_ncu.log 2>&1; echo "smoke under ncu rc=$?"; tail -2 gpurun_out/c11_smoke_ncu.log | cut -c1-200
