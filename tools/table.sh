#!/bin/bash
# The throughput table north_star asks for, one JSON line per (workload, GPU count):
#   tools/table.sh N [tag]      -> gpurun_out/<tag>_<workload>_n<N>.json   (tag defaults to r2_table)
# N = 1 also runs the single-GPU-only rows (configs 1 and 2).  TABLE_ONLY="c3 c4" restricts the rows.  Run under gpurun (--gpus N); nothing here needs /root/reference.
N=${1:-1}; TAG=${2:-r2_table}
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
run() {   # name, bench args...
  local name=$1; shift
  if [ -n "$TABLE_ONLY" ] && ! echo " $TABLE_ONLY " | grep -q " $name "; then return; fi
  if [ "$N" = 1 ]; then python bench.py --gpus 1 "$@" > gpurun_out/${TAG}_${name}_n$N.json 2> gpurun_out/${TAG}_${name}_n$N.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N "$@" \
         > gpurun_out/${TAG}_${name}_n$N.json 2> gpurun_out/${TAG}_${name}_n$N.err; fi
  echo "$name n=$N: $(head -c 300 gpurun_out/${TAG}_${name}_n$N.json)"
}
if [ "$N" = 1 ]; then
  run c1 --config c1 --steps 3 --warmup 3 --no-cpu
  run c2 --config c2 --steps 3 --warmup 3 --no-cpu
fi
run c3 --config c3 --steps 3 --warmup 3 --no-cpu
run c4 --config c4 --steps 2 --warmup 3 --no-cpu
run spheres100k --config c5 --tris 0 --spheres 100000 --steps 2 --warmup 3 --no-cpu
run tris1M --config c5 --tris 1000000 --spheres 0 --steps 2 --warmup 3 --no-cpu
run c5 --config c5 --steps 1 --warmup 3 --no-cpu
# the same C4 frame through the C entry point that owns the whole box (one process, srt_render_group)
if [ -z "$TABLE_ONLY" ] || echo " $TABLE_ONLY " | grep -q " c4_group "; then
python bench.py --backend group --gpus $N --config c4 --steps 2 --warmup 3 > gpurun_out/${TAG}_c4_group_n$N.json 2> gpurun_out/${TAG}_c4_group_n$N.err
fi
echo "c4 group n=$N: $(head -c 300 gpurun_out/${TAG}_c4_group_n$N.json)"
