#!/bin/bash
# GN A/B: tracker GN timing (kernel_timing.py, 3 repetitions) for the default library and every A/B library present.
for rep in 1 2 3; do
for lib in default splatt3r_slam_b200/libb200slam_ab_*.so; do
  [ "$lib" == default ] || [ -f "$lib" ] || continue
  if [ "$lib" == default ]; then unset B200SLAM_LIB; name=default; else export B200SLAM_LIB=$PWD/$lib; name=$(basename $lib .so); fi
  echo -n "$name: "; timeout 300 python tools/kernel_timing.py --no-graph 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print({k:(round(v,1) if isinstance(v,float) else v) for k,v in d.items() if k.startswith('gn_tracker')})"
done; done
