#!/bin/bash
# Quick refine A/B: matching tests + refine timing for the default library and every A/B library present.
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-q}
for lib in default splatt3r_slam_b200/libb200slam_ab_*.so; do
  [ "$lib" == default ] || [ -f "$lib" ] || continue
  if [ "$lib" == default ]; then unset B200SLAM_LIB; name=default; else export B200SLAM_LIB=$PWD/$lib; name=$(basename $lib .so); fi
  [ -n "$SKIP_TESTS" ] || { timeout 600 python -m pytest tests/test_gpu_matching.py tests/test_gpu_match_fused.py -m gpu -x -q > $OUT/${TAG}_${name}_pytest.log 2>&1; echo "$name pytest rc=$?"; tail -1 $OUT/${TAG}_${name}_pytest.log; }
  timeout 300 python tools/refine_timing.py 2> $OUT/${TAG}_${name}.err | tee $OUT/${TAG}_${name}.json
done
