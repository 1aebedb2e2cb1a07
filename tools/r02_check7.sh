#!/bin/bash
# round-2 GPU check 7: kernel-cost bisect (library variants), SQL + no-alloc tests, the reworked bench at N=1
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
export PGQ_B200_BACKTRACE=1
for v in BASE NODONE NODYN DENSE ALL; do
  lib=tools/ab/libpgq_$v.so; opt=""
  [ $v = BASE ] && lib=duckdb-pgq.old_b200/libpgq_b200.so
  [ $v = NODYN ] && opt="dyn=0"
  [ $v = ALL ] && opt="dyn=0"
  echo "== $v $opt" >> gpurun_out/c7_ab.log
  PGQ_B200_LIB=$PWD/$lib timeout 300 python tools/trace_levels.py 24 512 $opt 2>&1 | grep "iter [2-6]\|level_kernel_ms" | sed -E 's/next:.*//; s/.*level_kernel_ms.: ([0-9.]+).*/total \1/' >> gpurun_out/c7_ab.log
done
cat gpurun_out/c7_ab.log
timeout 900 python -m pytest tests/test_sql_extension.py tests/test_zz_no_alloc.py -m gpu -q -p no:faulthandler > gpurun_out/c7_sql.log 2>&1
echo "sql+noalloc rc=$?" | tee -a gpurun_out/c7_sql.log
tail -15 gpurun_out/c7_sql.log | cut -c1-250
timeout 1500 python bench.py --steps 3 --warmup 3 > gpurun_out/c7_bench.json 2> gpurun_out/c7_bench.err
echo "bench rc=$?"
tail -5 gpurun_out/c7_bench.err | cut -c1-300
python - <<'PY'
import json
d=json.load(open('gpurun_out/c7_bench.json'))
print("GTEPS",round(d['value']),"ms",round(d['ms_per_step'],1),"e2e",round(d['e2e']['value']),"chunked",round(d['e2e_chunked']['value']))
print("roofline frac",d['roofline']['frac'],"ref",d['roofline']['reference_model']['frac'],"syncs",d['host_syncs_in_level_loops'])
print("parity",d['parity'])
print("csr",{k:d['csr_build'][k] for k in ('value','ms','parity')})
p=d['paths']; print("paths",{k:(p[k].get('pairs_per_s'),p[k].get('e2e_chunked',{}).get('pairs_per_s')) for k in ('shortestpath','iterativelength')},p.get('parity'),p.get('cheapest_path_length'))
print("c1",d['config1']['gpu']['pairs_per_s'],d['config1'].get('cpu_oracle_1thread',{}).get('parity'))
print("cpu",d.get('cpu_baseline'))
PY
