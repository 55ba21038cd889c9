#!/bin/bash
# final-build regression check of config 5 (one 4-spp frame, both tree forms) + a default bench line
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 900 python tools/sweep_c5.py 10000000 100000 4 compact,full > $O/aa_c5.log 2>&1; cat $O/aa_c5.log | tail -n 4
timeout 600 python bench.py > $O/aa_bench_c4.json 2> $O/aa_bench_c4.err; echo "bench rc=$?"; cut -c1-200 $O/aa_bench_c4.json
