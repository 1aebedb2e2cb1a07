#!/bin/bash
# round-2 GPU check 3: locate the multi-batch shortestpath crash (backtrace on stderr, -s so pytest does not swallow it)
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
export PGQ_B200_BACKTRACE=1
rm -f /tmp/pgq_b200_crash.txt
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -s -p no:faulthandler -k "test_lengths_all_degree_bins and general" > gpurun_out/c3_crash.log 2>&1
echo "crash-test rc=$?" >> gpurun_out/c3_crash.log
cp /tmp/pgq_b200_crash.txt gpurun_out/c3_crash_bt.txt 2>/dev/null
tail -40 gpurun_out/c3_crash.log
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -s -p no:faulthandler -k "test_shortestpath_bit_identical or test_batches_in_flight or test_level_history_budget" > gpurun_out/c3_paths.log 2>&1
echo "paths rc=$?" >> gpurun_out/c3_paths.log
tail -30 gpurun_out/c3_paths.log
cp /tmp/pgq_b200_crash.txt gpurun_out/c3_crash_bt2.txt 2>/dev/null
