#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; L=$O/y_shade3.log; : > $L
for sh in 2 3 2 3; do
  echo "== SRT_WF_SHADE=$sh" >> $L
  SRT_WF_SHADE=$sh timeout 300 python tools/wf_quick.py 200 2>&1 | grep v2 >> $L
done
echo "== parity, SRT_WF_SHADE=3" >> $L
SRT_WF_SHADE=3 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_world.py -m gpu -x -q 2>&1 | tail -n 4 >> $L
cat $L
