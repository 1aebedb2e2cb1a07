#!/bin/bash
# round-2 GPU check 6: explicit-node loop graphs (early template destroy vs leak), dyn A/B, full suite
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
export PGQ_B200_BACKTRACE=1
rm -f /tmp/pgq_b200_crash.txt
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -s -p no:faulthandler -k "general" > gpurun_out/c6_general_early.log 2>&1
rc=$?
echo "general early rc=$rc" | tee -a gpurun_out/c6_general_early.log
grep -n "fatal signal" -A 16 gpurun_out/c6_general_early.log | cut -c1-160 | head -24
tail -3 gpurun_out/c6_general_early.log | cut -c1-300
MODE=early
if [ $rc -ne 0 ]; then
  export PGQ_B200_GRAPH_TEMPLATE=leak
  MODE=leak
  timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -s -p no:faulthandler -k "general" > gpurun_out/c6_general_leak.log 2>&1
  echo "general leak rc=$?" | tee -a gpurun_out/c6_general_leak.log
  grep -n "fatal signal" -A 16 gpurun_out/c6_general_leak.log | cut -c1-160 | head -24
  tail -3 gpurun_out/c6_general_leak.log | cut -c1-300
fi
echo "MODE=$MODE"
for v in "" "dyn=0"; do
  echo "== $v" >> gpurun_out/c6_traces.log
  timeout 300 python tools/trace_levels.py 24 512 $v 2>&1 | grep "iter\|level_kernel_ms" | cut -c1-200 >> gpurun_out/c6_traces.log
done
grep -v "^{" gpurun_out/c6_traces.log
timeout 1800 python -m pytest tests -m gpu -q -p no:faulthandler > gpurun_out/c6_pytest.log 2>&1
echo "pytest ($MODE) rc=$?" | tee -a gpurun_out/c6_pytest.log
tail -25 gpurun_out/c6_pytest.log | cut -c1-300
cp /tmp/pgq_b200_crash.txt gpurun_out/c6_crash_bt.txt 2>/dev/null
