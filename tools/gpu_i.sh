#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; : > $O/i_ab.log
for lib in libsrt_b200.so ab/e1.so ab/e2.so ab/e12.so ab/e14.so libsrt_b200.so; do
  echo "== $lib" >> $O/i_ab.log
  SRT_B200_LIBRARY=$PWD/sexy-raytracer_b200/$lib timeout 300 python tools/wf_quick.py 200 2>&1 | grep v2 >> $O/i_ab.log
done
for wf in "16777216,1024,16" "4194304,1024,16" "8388608,1024,12" "8388608,1024,20"; do
  echo "== SRT_WF=$wf" >> $O/i_ab.log
  SRT_WF=$wf timeout 300 python tools/wf_quick.py 200 2>&1 | grep v2 >> $O/i_ab.log
done
cat $O/i_ab.log
