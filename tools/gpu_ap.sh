#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; L=$O/ap_pool.log; : > $L
timeout 1400 python tools/sweep_c5.py 10000000 100000 64 compact "8388608,768,16" "33554432,768,16" "67108864,768,16" "134217728,768,16" >> $L 2>&1
cat $L
