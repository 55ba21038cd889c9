#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; : > $O/m_l2.log
for pct in 0 30 60 100; do
  echo "== SRT_L2_PERSIST=$pct" >> $O/m_l2.log
  SRT_L2_PERSIST=$pct timeout 600 python tools/sweep_c5.py 10000000 100000 4 compact >> $O/m_l2.log 2>&1
done
timeout 600 python bench.py > $O/m_bench_c4.json 2> $O/m_bench_c4.err; echo "rc=$?" >> $O/m_bench_c4.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/m_launches_bench.csv \
   python bench.py --steps 2 --warmup 1 --no-cpu > $O/m_ncu_bench.log 2>&1
cat $O/m_l2.log; tail -n 2 $O/m_bench_c4.err; cut -c1-300 $O/m_bench_c4.json; wc -l $O/m_launches_bench.csv
