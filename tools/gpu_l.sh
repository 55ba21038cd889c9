#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/l_pytest.log 2>&1; echo "pytest rc=$?" >> $O/l_pytest.log
timeout 900 python tools/sweep_c5.py 10000000 100000 4 full,compact > $O/l_sweep.log 2>&1
tail -n 6 $O/l_pytest.log; cat $O/l_sweep.log
