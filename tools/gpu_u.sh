#!/bin/bash
# A/B: tree in shared memory (MODE 0) vs eight octant-sorted copies through L1/L2 (MODE 3)
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; L=$O/u_octant.log; : > $L
echo "== smem (default)" >> $L
timeout 300 python tools/wf_quick.py 200 2>&1 | grep v2 >> $L
for wf in "8388608,1024,16" "8388608,768,16" "8388608,512,16"; do
  echo "== MODE 3 (octant copies), SRT_WF=$wf" >> $L
  SRT_FORCE_GLOBAL_TREE=1 SRT_WF_TREE=octant SRT_WF=$wf timeout 300 python tools/wf_quick.py 200 2>&1 | grep v2 >> $L
done
echo "== parity with octant copies" >> $L
SRT_FORCE_GLOBAL_TREE=1 SRT_WF_TREE=octant timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -n 5 >> $L
cat $L
