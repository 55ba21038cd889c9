#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_wf_extend -s 6 -c 1 -o $O/q_extend -f \
   python tools/prof_one.py c3_masterchief 1280 720 40 4 2 > $O/q_ncu.log 2>&1
ncu -i $O/q_extend.ncu-rep --page raw --csv > $O/q_extend_raw.csv 2>/dev/null
ncu -i $O/q_extend.ncu-rep --page source --csv --print-source cuda,sass > $O/q_extend_src.csv 2>/dev/null
tail -n 2 $O/q_ncu.log
