#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,lts__t_sectors_srcunit_tex_op_read.sum,smsp__issue_active.avg.pct_of_peak_sustained_active \
  --clock-control none -k regex:k_wf_extend -s 2 -c 1 --csv --log-file $O/aq_ncu_c5_64Mi.csv python tools/prof_c5.py 10000000 100000 64 > $O/aq_ncu.log 2>&1
echo "rc=$?"; tail -n 8 $O/aq_ncu_c5_64Mi.csv | cut -d, -f5,13,14,15 
