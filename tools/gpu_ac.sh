#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; L=$O/ac_sort.log; : > $L
for so in 0 1; do
  echo "== SRT_WF_SORT=$so, 16 spp" >> $L
  SRT_WF_SORT=$so timeout 900 python tools/sweep_c5.py 10000000 100000 16 compact,full >> $L 2>&1
done
SRT_WF_SORT=1 timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,lts__t_sectors_srcunit_tex_op_read.sum,l1tex__t_sector_hit_rate.pct,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio \
  --clock-control none -k regex:k_wf_extend -s 2 -c 2 --csv --log-file $O/ac_ncu_sorted.csv python tools/prof_c5.py 10000000 100000 8 > $O/ac_ncu.log 2>&1
cat $L; cat $O/ac_ncu_sorted.csv | tail -n 20
