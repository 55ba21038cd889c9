#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/ak_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 2 $O/ak_pytest.log
timeout 1500 python bench.py --config c5 --steps 2 --warmup 3 > $O/ak_bench_c5.json 2> $O/ak_bench_c5.err; echo "c5 rc=$?"; cut -c1-200 $O/ak_bench_c5.json; tail -n 2 $O/ak_bench_c5.err
