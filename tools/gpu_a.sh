#!/bin/bash
# one GPU session: parity tests, smoke, C4 + C5 bench lines
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/a_smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/a_pytest.log 2>&1; echo "pytest rc=$?" >> $O/a_pytest.log
timeout 300 python __graft_entry__.py smoke > $O/a_smoke.log 2>&1; echo "smoke rc=$?" >> $O/a_smoke.log
timeout 600 python bench.py > $O/a_bench_c4.json 2> $O/a_bench_c4.err; echo "c4 rc=$?" >> $O/a_bench_c4.err
timeout 600 python bench.py --config c5 --tris 1000000 --steps 2 --warmup 3 > $O/a_bench_c5_1m.json 2> $O/a_bench_c5_1m.err; echo "rc=$?" >> $O/a_bench_c5_1m.err
timeout 900 python bench.py --config c5 --steps 2 --warmup 3 > $O/a_bench_c5.json 2> $O/a_bench_c5.err; echo "rc=$?" >> $O/a_bench_c5.err
tail -3 $O/a_pytest.log $O/a_smoke.log $O/a_bench_c4.err $O/a_bench_c5_1m.err $O/a_bench_c5.err
cat $O/a_bench_c4.json $O/a_bench_c5_1m.json $O/a_bench_c5.json
