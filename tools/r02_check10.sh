#!/bin/bash
# round-2 GPU check 10: FILTER-tied dense/sparse kernels vs round 1, full suite, L2-window evidence, ncu launch lists
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for v in R01 BASE; do
  lib=tools/ab/libpgq_$v.so
  [ $v = BASE ] && lib=duckdb-pgq.old_b200/libpgq_b200.so
  echo "== $v" >> gpurun_out/c10_ab.log
  PGQ_B200_LIB=$PWD/$lib timeout 300 python tools/trace_levels.py 24 512 2>&1 | grep "iter [1-8]\|level_kernel_ms\|rror" | sed -E 's/next:.*//; s/.*level_kernel_ms.: ([0-9.]+).*/total \1/' >> gpurun_out/c10_ab.log
done
echo "== BASE l2_persist=1" >> gpurun_out/c10_ab.log
timeout 300 python tools/trace_levels.py 24 512 l2_persist=1 2>&1 | grep "iter [1-8]\|level_kernel_ms\|rror" | sed -E 's/next:.*//; s/.*level_kernel_ms.: ([0-9.]+).*/total \1/' >> gpurun_out/c10_ab.log
cat gpurun_out/c10_ab.log
timeout 1800 python -m pytest tests -m gpu -x -q -p no:faulthandler > gpurun_out/c10_pytest.log 2>&1
echo "pytest rc=$?" | tee -a gpurun_out/c10_pytest.log
tail -6 gpurun_out/c10_pytest.log | cut -c1-250
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
timeout 900 ncu --metrics $M --clock-control none --csv --log-file gpurun_out/c10_launches.csv python bench.py --pairs 2048 --steps 1 --warmup 1 --no-cpu-baseline --no-csr --no-paths > gpurun_out/c10_ncu_bench.log 2>&1
echo "ncu bench rc=$?"
timeout 600 ncu --metrics $M --clock-control none --csv --log-file gpurun_out/c10_csr_launches.csv python tools/csr_only.py 22 1 noprep > gpurun_out/c10_ncu_csr.log 2>&1
echo "ncu csr rc=$?"
timeout 1500 python bench.py --steps 5 --warmup 3 > gpurun_out/c10_bench.json 2> gpurun_out/c10_bench.err
echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/c10_bench.json'))
p=d['paths']
print("BENCH", round(d['value']), round(d['ms_per_step'],1), 'e2e', round(d['e2e']['value']), 'chunked', round(d['e2e_chunked']['value']), 'frac', round(d['roofline']['frac'],3),
      'C4 sp', round(p['shortestpath']['pairs_per_s']), 'len', round(p['iterativelength']['pairs_per_s']), 'c1', round(d['config1']['gpu']['pairs_per_s']), d['parity'])
PY
