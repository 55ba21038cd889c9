#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; : > $O/j_ab.log
timeout 900 python -m pytest tests -m gpu -x -q > $O/j_pytest.log 2>&1; echo "pytest rc=$?" >> $O/j_pytest.log
for sh in 1 2 1 2; do
  echo "== SRT_WF_SHADE=$sh" >> $O/j_ab.log
  SRT_WF_SHADE=$sh timeout 300 python tools/wf_quick.py 200 2>&1 | grep v2 >> $O/j_ab.log
done
tail -n 4 $O/j_pytest.log; cat $O/j_ab.log
