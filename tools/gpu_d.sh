#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/d_pytest.log 2>&1; echo "pytest rc=$?" >> $O/d_pytest.log
timeout 900 python tools/sweep_c5.py 10000000 100000 4 full,compact "8388608,1024,16" "8388608,1024,24" "8388608,1024,8" > $O/d_sweep.log 2>&1
timeout 600 python tools/sweep_c5.py 1000000 100000 8 full,compact >> $O/d_sweep.log 2>&1
timeout 1200 ncu --set full --import-source on --clock-control none -k regex:k_wf_extend -s 3 -c 1 -o $O/d_c5_extend_compact -f \
   python tools/prof_c5.py 10000000 100000 2 > $O/d_c5_ncu_full.log 2>&1
ncu -i $O/d_c5_extend_compact.ncu-rep --page raw --csv > $O/d_c5_extend_compact_raw.csv 2>/dev/null
tail -n 5 $O/d_pytest.log; cat $O/d_sweep.log; tail -n 3 $O/d_c5_ncu_full.log
