#!/bin/bash
# A/B session for refine_matches builds: full GPU tests on the default library, then matching tests +
# refine timing per A/B library (splatt3r_slam_b200/libb200slam_ab_*.so from tools/build_ab.sh).
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-ab}
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/${TAG}_pytest.log
timeout 300 python tools/refine_timing.py 2> $OUT/${TAG}_default.err | tee $OUT/${TAG}_default.json
for lib in splatt3r_slam_b200/libb200slam_ab_*.so; do
  name=$(basename $lib .so)
  B200SLAM_LIB=$PWD/$lib timeout 600 python -m pytest tests/test_gpu_matching.py tests/test_gpu_match_fused.py -m gpu -x -q > $OUT/${TAG}_${name}_pytest.log 2>&1; echo "$name pytest rc=$?"; tail -1 $OUT/${TAG}_${name}_pytest.log
  B200SLAM_LIB=$PWD/$lib timeout 300 python tools/refine_timing.py 2> $OUT/${TAG}_${name}.err | tee $OUT/${TAG}_${name}.json
done
