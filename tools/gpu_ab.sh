#!/bin/bash
# A/B timing session (tuning aid): bash tools/gpu_ab.sh <tag>
TAG=${1:-ab}; OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/${TAG}_pytest.log
timeout 300 python tools/kernel_timing.py > $OUT/${TAG}_default.json 2> $OUT/${TAG}_default.err; cat $OUT/${TAG}_default.json
B200SLAM_IP_V1=1 timeout 300 python tools/kernel_timing.py > $OUT/${TAG}_ipv1.json 2>&1; cat $OUT/${TAG}_ipv1.json
for t in 148 444 592 1184; do
B200SLAM_GN_TARGET_CTAS=$t timeout 300 python tools/kernel_timing.py > $OUT/${TAG}_ctas$t.json 2>&1; cat $OUT/${TAG}_ctas$t.json
done
