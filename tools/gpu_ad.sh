#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; L=$O/ad_runs.log; : > $L
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "hbm or compact" 2>&1 | tail -n 4 >> $L
echo "== run 4096 (default build), sort 1, 16 spp" >> $L
SRT_WF_SORT=1 timeout 900 python tools/sweep_c5.py 10000000 100000 16 compact,full >> $L 2>&1
for r in 1024 16384; do
  echo "== run $r, sort 1, 16 spp" >> $L
  SRT_B200_LIBRARY=$PWD/gpurun_ab_run$r.so SRT_WF_SORT=1 timeout 900 python tools/sweep_c5.py 10000000 100000 16 compact,full >> $L 2>&1
done
echo "== run 4096, sort 0" >> $L
SRT_WF_SORT=0 timeout 900 python tools/sweep_c5.py 10000000 100000 16 compact >> $L 2>&1
cat $L
