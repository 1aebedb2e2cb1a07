#!/bin/bash
# round-2 GPU check 20: occupancy A/B for the gather kernels (minimum blocks per SM)
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for v in BASE L12W16 L12 W16S16 W16 BASE; do
  lib=tools/ab/libpgq_$v.so; [ $v = BASE ] && lib=duckdb-pgq.old_b200/libpgq_b200.so
  echo "== $v"
  PGQ_TRACE_HOSTLOOP=1 PGQ_B200_LIB=$PWD/$lib timeout 300 python tools/trace_levels.py 24 512 2>&1 | grep "iter [2-6]\|level_kernel_ms" | sed -E 's/next:.*//; s/.*level_kernel_ms.: ([0-9.]+).*/total \1/'
done
for v in BASE L12W16 L12 W16S16 W16 BASE; do
  lib=tools/ab/libpgq_$v.so; [ $v = BASE ] && lib=duckdb-pgq.old_b200/libpgq_b200.so
  PGQ_B200_LIB=$PWD/$lib timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-csr > gpurun_out/c20_bench_$v.json 2> gpurun_out/c20_bench_$v.err
  python - <<PY
import json
d=json.load(open('gpurun_out/c20_bench_$v.json'))
p=d['paths']
print("BENCH $v", round(d['value']), round(d['ms_per_step'],1), 'C4 sp', round(p['shortestpath']['pairs_per_s']), 'len', round(p['iterativelength']['pairs_per_s']))
PY
done
