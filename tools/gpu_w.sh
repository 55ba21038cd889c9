#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/w_pytest.log 2>&1
tail -n 25 $O/w_pytest.log
echo "== quick timing (regression check)"; timeout 300 python tools/wf_quick.py 200 2>&1 | grep v2
