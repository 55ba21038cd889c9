#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; L=$O/ag_l1.log; : > $L
echo "== default carveout" >> $L
timeout 900 python tools/sweep_c5.py 10000000 100000 16 compact >> $L 2>&1
for c in 0 25 100; do
  echo "== SRT_WF_L1=$c (preferred shared-memory carveout, percent)" >> $L
  SRT_WF_L1=$c timeout 900 python tools/sweep_c5.py 10000000 100000 16 compact >> $L 2>&1
done
cat $L
