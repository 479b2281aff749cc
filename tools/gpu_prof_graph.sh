#!/bin/bash
TAG=${1:-r01b}; OUT=gpurun_out; mkdir -p $OUT
timeout 600 python tools/prof_graph.py > $OUT/${TAG}_plain_graph.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'build_kernel|solve_retract|finalize_kernel|assemble_kernel' -s 4 -c 4 \
    -o $OUT/${TAG}_graph_prof -f python tools/prof_graph.py > $OUT/${TAG}_ncu_graph.log 2>&1
echo "ncu rc=$?"; tail -2 $OUT/${TAG}_ncu_graph.log
