#!/bin/bash
# round-2 GPU check 34: SQL tests after the extension gained pgq_rewrite()
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
timeout 900 python -m pytest tests/test_pgq_rewrite.py tests/test_sql_extension.py tests/test_zz_no_alloc.py -q -p no:faulthandler 2>&1 | tail -3
