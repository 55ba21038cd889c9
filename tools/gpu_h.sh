#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
for lib in libsrt_b200.so ab/libsrt_mb4.so ab/libsrt_mb5.so libsrt_b200.so; do
  echo "== $lib" >> $O/h_shade_ab.log
  SRT_B200_LIBRARY=$PWD/sexy-raytracer_b200/$lib timeout 300 python tools/wf_quick.py 200 >> $O/h_shade_ab.log 2>&1
done
timeout 1200 ncu --set full --import-source on --clock-control none -k regex:k_wf_extend -s 2 -c 1 -o $O/h_c5_extend_steady -f \
   python tools/prof_c5.py 10000000 100000 8 > $O/h_c5_ncu_full.log 2>&1
ncu -i $O/h_c5_extend_steady.ncu-rep --page raw --csv > $O/h_c5_extend_steady_raw.csv 2>/dev/null
cat $O/h_shade_ab.log; tail -n 3 $O/h_c5_ncu_full.log
