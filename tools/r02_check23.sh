#!/bin/bash
# round-2 GPU check 23: radix scatter ranking with MATCH.ANY against the per-bit ballots
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for lib in duckdb-pgq.old_b200/libpgq_b200.so tools/ab/libpgq_RWMATCH.so duckdb-pgq.old_b200/libpgq_b200.so tools/ab/libpgq_RWMATCH.so; do
  echo "== $lib"
  PGQ_B200_LIB=$PWD/$lib timeout 300 python tools/csr_only.py 22 6 noprep 2>&1 | tail -4 | tr '\n' ' '; echo
  PGQ_B200_LIB=$PWD/$lib timeout 300 python tools/csr_only.py 24 4 noprep 2>&1 | tail -3 | tr '\n' ' '; echo
done
PGQ_B200_LIB=$PWD/tools/ab/libpgq_RWMATCH.so timeout 900 python -m pytest tests/test_gpu_parity.py -q -x -p no:faulthandler -k "csr" 2>&1 | tail -3
