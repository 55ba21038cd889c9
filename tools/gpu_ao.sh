#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; L=$O/ao_pool.log; : > $L
timeout 1200 python tools/sweep_c5.py 10000000 100000 32 compact "8388608,768,16" "16777216,768,16" "33554432,768,16" "4194304,768,16" >> $L 2>&1
cat $L
