#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; : > $O/t_mode1.log
echo "== smem (default)" >> $O/t_mode1.log
timeout 300 python tools/wf_quick.py 200 2>&1 | grep v2 >> $O/t_mode1.log
for wf in "8388608,1024,16" "8388608,768,16" "8388608,512,16"; do
  echo "== MODE 1 (32-byte nodes through L1/L2), SRT_WF=$wf" >> $O/t_mode1.log
  SRT_FORCE_GLOBAL_TREE=1 SRT_WF_TREE=full SRT_WF=$wf timeout 300 python tools/wf_quick.py 200 2>&1 | grep v2 >> $O/t_mode1.log
done
cat $O/t_mode1.log
