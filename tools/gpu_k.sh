#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_wf_shade2 -s 6 -c 1 -o $O/k_shade2 -f \
   python tools/prof_one.py c3_masterchief 1280 720 40 4 2 > $O/k_ncu.log 2>&1
ncu -i $O/k_shade2.ncu-rep --page raw --csv > $O/k_shade2_raw.csv 2>/dev/null
ncu -i $O/k_shade2.ncu-rep --page source --csv --print-source cuda,sass > $O/k_shade2_src.csv 2>/dev/null
tail -n 3 $O/k_ncu.log
