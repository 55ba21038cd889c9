#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; L=$O/ah_threads.log; : > $L
for cfg in "256 1" "256 2" "512 1" "768 1"; do
  set -- $cfg
  echo "== $1 threads x $2 CTA per SM" >> $L
  SRT_WF_CTAS_PER_SM=$2 timeout 900 python tools/sweep_c5.py 10000000 100000 16 compact "8388608,$1,16" >> $L 2>&1
done
cat $L
