#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 600 python tools/prof_c5.py 10000000 100000 8 > $O/an_prof.log 2>&1; rc=$?; tail -n 2 $O/an_prof.log
if [ $rc -eq 0 ]; then
  timeout 1200 ncu --set full --import-source on --clock-control none -k regex:k_wf_extend -s 2 -c 1 -o $O/an_c5_extend -f \
    python tools/prof_c5.py 10000000 100000 8 > $O/an_ncu.log 2>&1; echo "ncu rc=$?"
  ncu -i $O/an_c5_extend.ncu-rep --page raw --csv > $O/an_c5_extend_raw.csv 2>/dev/null
  ncu -i $O/an_c5_extend.ncu-rep --page source --csv --print-source cuda,sass > $O/an_c5_extend_src.csv 2>/dev/null
  ls -la $O/an_*
fi
