#!/bin/bash
# Multi-GPU session: bash tools/gpu_multi.sh <tag> <ngpus>
TAG=${1:-mg}; N=${2:-2}; OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 > $OUT/${TAG}_tracker_n$N.json 2> $OUT/${TAG}_tracker_n$N.err; echo "tracker rc=$?"; cat $OUT/${TAG}_tracker_n$N.json | cut -c1-400
timeout 600 $TR bench.py --gpus $N --workload graph --steps 10 --warmup 3 > $OUT/${TAG}_graph_n$N.json 2> $OUT/${TAG}_graph_n$N.err; echo "graph rc=$?"; cat $OUT/${TAG}_graph_n$N.json | cut -c1-600
timeout 600 $TR bench.py --gpus $N --impl reference --steps 2 --warmup 1 > $OUT/${TAG}_ref_n$N.json 2> $OUT/${TAG}_ref_n$N.err; echo "ref rc=$?"; cat $OUT/${TAG}_ref_n$N.json | cut -c1-200
tail -n 3 $OUT/${TAG}_tracker_n$N.err; tail -n 3 $OUT/${TAG}_graph_n$N.err; exit 0
