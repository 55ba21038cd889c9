#!/bin/bash
# usage: gpu_n.sh "<list of N>"   - bench.py at each N (C4), no CPU leg
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/n_smi.txt
for N in $1; do
  if [ "$N" = "1" ]; then
    timeout 600 python bench.py --gpus 1 --steps 3 --warmup 3 --no-cpu > $O/n_bench_c4_n1.json 2> $O/n_bench_c4_n1.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
      bench.py --gpus $N --steps 3 --warmup 3 > $O/n_bench_c4_n$N.json 2> $O/n_bench_c4_n$N.err
  fi
  echo "N=$N rc=$?"; tail -n 2 $O/n_bench_c4_n$N.err; cat $O/n_bench_c4_n$N.json | cut -c1-400
done
