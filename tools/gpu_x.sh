#!/bin/bash
# Round-end evidence on ONE B200 with the final build: GPU tests, smoke, both bench arms, ncu launch list, one full capture of
# each wavefront kernel (each ncu pass only after its own command exited 0 without ncu).
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/x_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 2 $O/x_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/x_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 $O/x_smoke.log
timeout 900 python bench.py --impl reference > $O/x_bench_ref.json 2> $O/x_bench_ref.err; echo "ref rc=$?"; cut -c1-300 $O/x_bench_ref.json
timeout 900 python bench.py > $O/x_bench_c4.json 2> $O/x_bench_c4.err; echo "bench rc=$?"; cut -c1-400 $O/x_bench_c4.json
timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu > $O/x_b21.json 2> $O/x_b21.err; rc=$?; echo "bench 2/1 rc=$rc"
if [ $rc -eq 0 ]; then
  timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $O/x_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu > $O/x_ncu_launch.log 2>&1; echo "ncu launches rc=$?"; wc -l $O/x_launches.csv
fi
timeout 300 python tools/prof_one.py c3_masterchief 1280 720 40 4 2 > $O/x_prof_one.log 2>&1; rc=$?; echo "prof_one rc=$rc"
if [ $rc -eq 0 ]; then
  for k in k_wf_extend k_wf_shade2; do
    timeout 900 ncu --set full --import-source on --clock-control none -k regex:$k -s 6 -c 1 -o $O/x_$k -f \
      python tools/prof_one.py c3_masterchief 1280 720 40 4 2 > $O/x_ncu_$k.log 2>&1; echo "ncu $k rc=$?"
    ncu -i $O/x_$k.ncu-rep --page raw --csv > $O/x_${k}_raw.csv 2>/dev/null
  done
fi
ls -la $O | tail -n 20
