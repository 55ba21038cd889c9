#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; L=$O/ae_keys.log; : > $L
for k in 0 1 2 3 4; do
  echo "== SRT_WF_KEY=$k, 16 spp" >> $L
  SRT_WF_KEY=$k timeout 900 python tools/sweep_c5.py 10000000 100000 16 compact >> $L 2>&1
done
cat $L
