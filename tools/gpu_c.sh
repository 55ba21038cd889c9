#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 1500 python tools/sweep_c5.py 10000000 100000 4 0,1,2,3,5,7,9,11,13,15 "8388608,1024,16" > $O/c_sweep.log 2>&1
timeout 600 python tools/sweep_c5.py 10000000 100000 4 0,3 "8388608,1024,8" "8388608,1024,24" "8388608,1024,30" "2097152,1024,16" >> $O/c_sweep.log 2>&1
cat $O/c_sweep.log
