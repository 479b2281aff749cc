#!/bin/bash
# Strong-scaling run of the edge-sharded GN: bash tools/gpu_scale.sh <tag> <max_gpus> <kfs> <edges>
TAG=${1:-sc}; MAXN=${2:-8}; KFS=${3:-32}; EDGES=${4:-896}; OUT=gpurun_out; mkdir -p $OUT
for N in 1 2 4 8; do
  [ $N -gt $MAXN ] && break
  if [ $N -eq 1 ]; then CMD="python"; else CMD="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N"; fi
  timeout 900 $CMD bench.py --gpus $N --workload graph --kfs $KFS --edges $EDGES --steps 5 --warmup 3 > $OUT/${TAG}_graph_k${KFS}_e${EDGES}_n$N.json 2> $OUT/${TAG}_n$N.err
  echo "N=$N rc=$?"; cut -c1-130 $OUT/${TAG}_graph_k${KFS}_e${EDGES}_n$N.json; tail -n 2 $OUT/${TAG}_n$N.err
done
if [ $MAXN -ge 8 ]; then
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus 8 --steps 20 --warmup 5 > $OUT/${TAG}_tracker_n8.json 2> $OUT/${TAG}_tracker_n8.err; echo "tracker8 rc=$?"; cut -c1-130 $OUT/${TAG}_tracker_n8.json
fi
exit 0
