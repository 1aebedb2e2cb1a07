#!/bin/bash
# round-2 GPU check 32: fractional DOUBLE weights (8-byte rows): batches in flight
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
for s in 1 2 3 4; do
  PROBE_DOUBLE=1 timeout 300 python tools/sssp_probe.py 1024 sssp_slots=$s 2>&1 | tail -1
done
timeout 300 python tools/sssp_probe.py 1024 sssp_slots=4 sssp_narrow=0 2>&1 | tail -1
timeout 300 python tools/sssp_probe.py 1024 sssp_slots=2 sssp_narrow=0 2>&1 | tail -1
timeout 300 python tools/sssp_probe.py 1024 sssp_slots=2 2>&1 | tail -1
timeout 300 python tools/sssp_probe.py 1024 sssp_slots=3 2>&1 | tail -1
