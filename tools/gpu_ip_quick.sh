#!/bin/bash
# iter_proj A/B timing (kernel_timing.py --no-graph) for the default library and every A/B library present.
for lib in default splatt3r_slam_b200/libb200slam_ab_*.so; do
  [ "$lib" == default ] || [ -f "$lib" ] || continue
  if [ "$lib" == default ]; then unset B200SLAM_LIB; name=default; else export B200SLAM_LIB=$PWD/$lib; name=$(basename $lib .so); fi
  echo -n "$name: "; timeout 300 python tools/kernel_timing.py --no-graph 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print({k:(round(v,2) if isinstance(v,float) else v) for k,v in d.items() if k in ('iter_proj_us','match_fused_us','refine_us','gn_tracker_10it_us')})"
done
