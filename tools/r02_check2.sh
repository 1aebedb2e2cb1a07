#!/bin/bash
# round-2 GPU check 2: locate the shortestpath crash, full tests, per-level traces of option variants, ncu list
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
export PGQ_B200_BACKTRACE=1
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -k "test_lengths_all_degree_bins and general" > gpurun_out/c2_crash.log 2>&1
echo "crash-test rc=$?" >> gpurun_out/c2_crash.log
grep -n "pgq_b200\]\|libpgq\|passed\|failed" gpurun_out/c2_crash.log | head -40
timeout 1200 python -m pytest tests -m gpu -q -n 2 > gpurun_out/c2_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c2_pytest.log
tail -15 gpurun_out/c2_pytest.log
for v in "" "dense_store=0" "done_bitmap=2" "done_bitmap=0" "dense_store=0 done_bitmap=2"; do
  echo "== $v" >> gpurun_out/c2_traces.log
  timeout 300 python tools/trace_levels.py 24 512 $v 2>&1 | grep "iter\|level_kernel_ms" >> gpurun_out/c2_traces.log
done
cat gpurun_out/c2_traces.log
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/c2_ncu_one_batch.csv python tools/trace_levels.py 24 512 > gpurun_out/c2_ncu.log 2>&1
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-csr > gpurun_out/c2_bench.json 2> gpurun_out/c2_bench.err
echo "bench rc=$?"
python -c "
import json;d=json.load(open('gpurun_out/c2_bench.json'));print(d['value'],d['ms_per_step'],d['e2e']['value'],d.get('paths'))"
