#!/bin/bash
# refine: band-count sweep on the default library + ncu capture of the default library.
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-sw}
for nb in 24 26 27 28 29 30 32; do
  echo -n "bands=$nb "; B200SLAM_REFINE_BANDS=$nb timeout 300 python tools/refine_timing.py 2>/dev/null | cut -c1-80
done
timeout 300 python tools/prof_kernels.py > $OUT/${TAG}_plain_prof.log 2>&1 || { echo "plain run failed"; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:refine_tile_kernel -s 2 -c 1 \
    -o $OUT/${TAG}_refine -f python tools/prof_kernels.py > $OUT/${TAG}_ncu_refine.log 2>&1; echo "ncu rc=$?"
