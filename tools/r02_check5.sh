#!/bin/bash
# round-2 GPU check 5: the silent crash of the full suite at test_lengths_all_degree_bins[general-256-bins0]
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
export PGQ_B200_BACKTRACE=1
which gdb valgrind > gpurun_out/c5_tools.txt 2>&1
rm -f /tmp/pgq_b200_crash.txt
ulimit -c 0
MALLOC_CHECK_=3 timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -s -p no:faulthandler -k "general" > gpurun_out/c5_full_general.log 2>&1
echo "general rc=$?" | tee -a gpurun_out/c5_full_general.log
cp /tmp/pgq_b200_crash.txt gpurun_out/c5_bt_general.txt 2>/dev/null
tail -30 gpurun_out/c5_full_general.log | cut -c1-300
rm -f /tmp/pgq_b200_crash.txt
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -s -p no:faulthandler -k "general" > gpurun_out/c5_full_general2.log 2>&1
echo "general (no malloc check) rc=$?" | tee -a gpurun_out/c5_full_general2.log
cp /tmp/pgq_b200_crash.txt gpurun_out/c5_bt_general2.txt 2>/dev/null
tail -30 gpurun_out/c5_full_general2.log | cut -c1-300
