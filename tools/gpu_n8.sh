#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29544"
timeout 600 python bench.py --gpus 1 --steps 3 --warmup 3 --no-cpu > $O/p_c4_n1.json 2> $O/p_c4_n1.err
for N in 2 4 8; do timeout 600 $TR --nproc-per-node $N bench.py --gpus $N --steps 3 --warmup 3 > $O/p_c4_n$N.json 2> $O/p_c4_n$N.err; done
timeout 900 $TR --nproc-per-node 8 bench.py --gpus 8 --config c5 --steps 2 --warmup 3 > $O/p_c5_n8.json 2> $O/p_c5_n8.err
for f in p_c4_n1 p_c4_n2 p_c4_n4 p_c4_n8 p_c5_n8; do tail -n 1 $O/$f.err | cut -c1-200; cut -c1-260 $O/$f.json; done
