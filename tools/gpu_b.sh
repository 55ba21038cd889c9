#!/bin/bash
# GPU session b: new host-side GPU test, config-5 ncu launch list + full-set capture of k_wf_extend (nodes in HBM)
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_world.py -m gpu -x -q > $O/b_pytest.log 2>&1; echo "pytest rc=$?" >> $O/b_pytest.log
timeout 600 python tools/prof_c5.py 10000000 100000 4 > $O/b_c5_plain.log 2>&1; echo "rc=$?" >> $O/b_c5_plain.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/b_launches_c5.csv \
   python tools/prof_c5.py 10000000 100000 2 > $O/b_c5_ncu_list.log 2>&1
timeout 1200 ncu --set full --import-source on --clock-control none -k regex:k_wf_extend -s 3 -c 1 -o $O/b_c5_extend -f \
   python tools/prof_c5.py 10000000 100000 2 > $O/b_c5_ncu_full.log 2>&1
ncu -i $O/b_c5_extend.ncu-rep --page raw --csv > $O/b_c5_extend_raw.csv 2>/dev/null
tail -n 4 $O/b_pytest.log $O/b_c5_plain.log $O/b_c5_ncu_list.log $O/b_c5_ncu_full.log
