#!/bin/bash
# iter_proj A/B: occupancy targets of the single-tap-set kernel.  Bit-exactness first, then timings.
TAG=${1:-r02i}; OUT=gpurun_out; mkdir -p $OUT
for L in "" ip4 ip8; do
  if [ -z "$L" ]; then export -n B200SLAM_LIB; unset B200SLAM_LIB; NAME=default; else export B200SLAM_LIB=$PWD/splatt3r_slam_b200/libb200slam_ab_$L.so; NAME=$L; fi
  timeout 900 python -m pytest tests/test_gpu_matching.py tests/test_gpu_match_fused.py -x -q > $OUT/${TAG}_pytest_$NAME.log 2>&1; echo "$NAME pytest rc=$?"; tail -1 $OUT/${TAG}_pytest_$NAME.log
  timeout 300 python tools/kernel_timing.py > $OUT/${TAG}_timing_$NAME.json 2> $OUT/${TAG}_timing_$NAME.err
  python -c "
import json; d=json.load(open('$OUT/${TAG}_timing_$NAME.json')); print('$NAME', {k: round(v,1) for k,v in d.items() if isinstance(v,(int,float))})"
done
