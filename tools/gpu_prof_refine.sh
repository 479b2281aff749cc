#!/bin/bash
# ncu --set full capture of refine_tile_kernel for the default library and every A/B library present.
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-r01g}
timeout 300 python tools/prof_kernels.py > $OUT/${TAG}_plain_prof.log 2>&1 || { echo "plain run failed"; tail -5 $OUT/${TAG}_plain_prof.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:refine_tile_kernel -s 2 -c 1 \
    -o $OUT/${TAG}_refine -f python tools/prof_kernels.py > $OUT/${TAG}_ncu_refine.log 2>&1; echo "ncu rc=$?"
for lib in splatt3r_slam_b200/libb200slam_ab_*.so; do
  [ -f $lib ] || continue
  name=$(basename $lib .so)
  B200SLAM_LIB=$PWD/$lib timeout 600 ncu --set full --clock-control none --import-source on -k regex:refine_tile_kernel -s 2 -c 1 \
      -o $OUT/${TAG}_refine_${name} -f python tools/prof_kernels.py > $OUT/${TAG}_ncu_refine_${name}.log 2>&1; echo "ncu $name rc=$?"
done
ls -la $OUT | tail -8
