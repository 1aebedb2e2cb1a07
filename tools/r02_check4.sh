#!/bin/bash
# round-2 GPU check 4: no-alloc entry point alone, the former graph-destroy crash, full suite, traces, bench
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
export PGQ_B200_BACKTRACE=1
rm -f /tmp/pgq_b200_crash.txt
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/c4_noalloc_alone.log 2>&1
echo "noalloc alone rc=$?" | tee -a gpurun_out/c4_noalloc_alone.log
tail -15 gpurun_out/c4_noalloc_alone.log
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -s -p no:faulthandler -k "test_lengths_all_degree_bins and general" > gpurun_out/c4_crash.log 2>&1
echo "crash-test rc=$?" | tee -a gpurun_out/c4_crash.log
tail -5 gpurun_out/c4_crash.log
cp /tmp/pgq_b200_crash.txt gpurun_out/c4_crash_bt.txt 2>/dev/null
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/c4_pytest.log 2>&1
echo "pytest rc=$?" | tee -a gpurun_out/c4_pytest.log
tail -25 gpurun_out/c4_pytest.log
cp /tmp/pgq_b200_crash.txt gpurun_out/c4_crash_bt2.txt 2>/dev/null
for v in "" "dense_store=0" "done_bitmap=0"; do
  echo "== $v" >> gpurun_out/c4_traces.log
  timeout 300 python tools/trace_levels.py 24 512 $v 2>&1 | grep "iter\|level_kernel_ms" >> gpurun_out/c4_traces.log
done
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/c4_ncu_one_batch.csv python tools/trace_levels.py 24 512 > gpurun_out/c4_ncu.log 2>&1
for opt in "" "--loop 0" "--streams 4" "--streams 16"; do
  timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-csr --no-paths $opt > gpurun_out/c4_bench.tmp 2>> gpurun_out/c4_bench.err
  echo "bench [$opt] rc=$?"
  python - <<PY
import json
try:
    d=json.load(open('gpurun_out/c4_bench.tmp'))
    print("BENCH [$opt]", round(d['value']), round(d['ms_per_step'],1), 'e2e', round(d['e2e']['value']), d['config'].get('batches_in_flight'))
    open('gpurun_out/c4_bench_lines.jsonl','a').write(json.dumps({"opt":"$opt","line":d})+"\n")
except Exception as ex:
    print("BENCH [$opt] failed", ex)
PY
done
