#!/bin/bash
# ncu --set full capture of the dense stage-3 kernel (one cooperative launch) on a 256-keyframe system
TAG=${1:-r01c}; OUT=gpurun_out; mkdir -p $OUT
timeout 300 python tools/dense_probe.py 256 2048 > $OUT/${TAG}_plain_dense.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'dense_solve_retract' -s 2 -c 1 \
    -o $OUT/${TAG}_dense_prof -f python tools/dense_probe.py 256 2048 > $OUT/${TAG}_ncu_dense.log 2>&1
echo "ncu rc=$?"; tail -2 $OUT/${TAG}_ncu_dense.log; cat $OUT/${TAG}_plain_dense.log
