#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; L=$O/z_prefetch.log; : > $L
for rep in 1 2; do
  echo "== prefetch off" >> $L
  SRT_B200_LIBRARY=$PWD/gpurun_ab_pf0.so timeout 300 python tools/wf_quick.py 200 2>&1 | grep v2 >> $L
  echo "== prefetch on (shade2)" >> $L
  timeout 300 python tools/wf_quick.py 200 2>&1 | grep v2 >> $L
done
cat $L
