#!/bin/bash
# round-2 GPU check 41: is the shortestpath jitter in the library? (C ABI, no engine)
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
timeout 300 python tools/sp_jitter.py 2>&1 | tail -3
timeout 300 python tools/sp_jitter.py streams=1 2>&1 | tail -3
timeout 300 python tools/sp_jitter.py small_path=0 2>&1 | tail -3
