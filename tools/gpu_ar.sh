#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 1
timeout 600 python bench.py > $O/ar_bench_c4.json 2> $O/ar_bench_c4.err; echo "bench rc=$?"; wc -l $O/ar_bench_c4.json; cut -c1-160 $O/ar_bench_c4.json
timeout 600 python bench.py --impl reference > $O/ar_bench_ref.json 2> $O/ar_bench_ref.err; echo "ref rc=$?"; cut -c1-120 $O/ar_bench_ref.json
