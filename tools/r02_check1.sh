#!/bin/bash
# round-2 GPU check 1: tests, per-level trace of one s24 batch (old vs new switches), a short bench
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/c1_gpus.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c1_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c1_pytest.log
tail -5 gpurun_out/c1_pytest.log
timeout 300 python tools/trace_levels.py 24 512 > gpurun_out/c1_trace_default.log 2>&1
timeout 300 python tools/trace_levels.py 24 512 done_bitmap=0 > gpurun_out/c1_trace_nodone.log 2>&1
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-csr --no-paths > gpurun_out/c1_bench_short.json 2> gpurun_out/c1_bench_short.err
echo "bench rc=$?"
tail -c 3000 gpurun_out/c1_bench_short.json
tail -5 gpurun_out/c1_bench_short.err
