#!/bin/bash
# round-2 GPU check 9: late view lookups (BASE) vs round 1 vs compile-time variants; C4 loop on/off
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for v in R01 BASE NODYN ND ALL BASE; do
  lib=tools/ab/libpgq_$v.so; opt=""; hl=""
  [ $v = BASE ] && lib=duckdb-pgq.old_b200/libpgq_b200.so
  [ $v = ALL ] && opt="dyn=0" && hl=1
  [ $v = NODYN ] && opt="dyn=0" && hl=1
  echo "== $v $opt" >> gpurun_out/c9_ab.log
  PGQ_TRACE_HOSTLOOP=$hl PGQ_B200_LIB=$PWD/$lib timeout 300 python tools/trace_levels.py 24 512 $opt 2>&1 | grep "iter [2-6]\|level_kernel_ms\|rror" | sed -E 's/next:.*//; s/.*level_kernel_ms.: ([0-9.]+).*/total \1/' >> gpurun_out/c9_ab.log
done
cat gpurun_out/c9_ab.log
for opt in "" "--loop 0"; do
  timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-csr $opt > gpurun_out/c9_bench.tmp 2>> gpurun_out/c9_bench.err
  python - <<PY
import json
d=json.load(open('gpurun_out/c9_bench.tmp'))
p=d['paths']
print("BENCH [$opt]", round(d['value']), round(d['ms_per_step'],1), 'e2e', round(d['e2e']['value']), 'chunked', round(d['e2e_chunked']['value']),
      'C4 sp', round(p['shortestpath']['pairs_per_s']), round(p['shortestpath']['e2e_chunked']['pairs_per_s']),
      'len', round(p['iterativelength']['pairs_per_s']), round(p['iterativelength']['e2e_chunked']['pairs_per_s']), 'c1', round(d['config1']['gpu']['pairs_per_s']))
open('gpurun_out/c9_bench_lines.jsonl','a').write(json.dumps({"opt":"$opt","line":d})+"\n")
PY
done
