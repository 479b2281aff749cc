#!/bin/bash
# config-3 solve timing (kernel_timing.py graph block) for the default library and every A/B library present.
for lib in default splatt3r_slam_b200/libb200slam_ab_*.so; do
  [ "$lib" == default ] || [ -f "$lib" ] || continue
  if [ "$lib" == default ]; then unset B200SLAM_LIB; name=default; else export B200SLAM_LIB=$PWD/$lib; name=$(basename $lib .so); fi
  echo -n "$name: "; timeout 300 python tools/kernel_timing.py 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['graph_N32_E128'])"
done
