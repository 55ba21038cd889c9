#!/bin/bash
# ray ordering (k_wf_keys + radix sort) on config 5: A/B at 4 spp, both tree forms; parity of the HBM-tree tests
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; L=$O/ab_sort.log; : > $L
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "hbm or compact" 2>&1 | tail -n 4 >> $L
for so in 0 1; do
  echo "== SRT_WF_SORT=$so" >> $L
  SRT_WF_SORT=$so timeout 900 python tools/sweep_c5.py 10000000 100000 4 compact,full >> $L 2>&1
done
echo "== 1 M triangles" >> $L
for so in 0 1; do
  echo "== SRT_WF_SORT=$so" >> $L
  SRT_WF_SORT=$so timeout 900 python tools/sweep_c5.py 1000000 10000 4 compact,full >> $L 2>&1
done
cat $L
