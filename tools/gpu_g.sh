#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/g_pytest.log 2>&1; echo "pytest rc=$?" >> $O/g_pytest.log
timeout 600 python bench.py > $O/g_bench_c4.json 2> $O/g_bench_c4.err; echo "rc=$?" >> $O/g_bench_c4.err
timeout 1500 python bench.py --config c5 --steps 2 --warmup 3 > $O/g_bench_c5.json 2> $O/g_bench_c5.err; echo "rc=$?" >> $O/g_bench_c5.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file $O/g_launches_c5.csv \
   python tools/prof_c5.py 10000000 100000 8 > $O/g_c5_ncu_list.log 2>&1
timeout 1200 ncu --set full --import-source on --clock-control none -k regex:k_wf_extend -s 20 -c 1 -o $O/g_c5_extend_steady -f \
   python tools/prof_c5.py 10000000 100000 8 > $O/g_c5_ncu_full.log 2>&1
ncu -i $O/g_c5_extend_steady.ncu-rep --page raw --csv > $O/g_c5_extend_steady_raw.csv 2>/dev/null
tail -n 3 $O/g_pytest.log $O/g_bench_c4.err $O/g_bench_c5.err $O/g_c5_ncu_full.log
cat $O/g_bench_c4.json $O/g_bench_c5.json
