#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 \
  bench.py --gpus 2 --config c5 --steps 2 --warmup 3 > $O/am_bench_c5_n2.json 2> $O/am_bench_c5_n2.err
echo "rc=$?"; tail -n 2 $O/am_bench_c5_n2.err; cut -c1-250 $O/am_bench_c5_n2.json
