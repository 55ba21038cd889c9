#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; L=$O/aj_noalloc.log; : > $L
echo "== lprims through L1 (old)" >> $L
SRT_B200_LIBRARY=$PWD/gpurun_ab_alloc.so timeout 900 python tools/sweep_c5.py 10000000 100000 16 compact >> $L 2>&1
echo "== lprims L1::no_allocate" >> $L
timeout 900 python tools/sweep_c5.py 10000000 100000 16 compact >> $L 2>&1
cat $L
