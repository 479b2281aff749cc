#!/bin/bash
# A/B session for GN builds: full GPU tests on the default library, then GN tests + tools/gn_timing-style numbers
# (kernel_timing.py) per A/B library.
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-gnab}
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $OUT/${TAG}_pytest.log
timeout 300 python tools/kernel_timing.py 2> $OUT/${TAG}_default.err | tee $OUT/${TAG}_default.json | cut -c1-900
for lib in splatt3r_slam_b200/libb200slam_ab_*.so; do
  [ -f $lib ] || continue
  name=$(basename $lib .so)
  B200SLAM_LIB=$PWD/$lib timeout 300 python tools/kernel_timing.py 2> $OUT/${TAG}_${name}.err | tee $OUT/${TAG}_${name}.json | cut -c1-900
done
