#!/bin/bash
# A/B timing session (tuning aid): bash tools/gpu_ab.sh <tag> [ENVVAR=value ...]
# Runs the GPU tests, then tools/kernel_timing.py once with the default settings and once per
# extra environment assignment given on the command line.
TAG=${1:-ab}; shift; OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $OUT/${TAG}_pytest.log
timeout 300 python tools/kernel_timing.py > $OUT/${TAG}_default.json 2> $OUT/${TAG}_default.err; cat $OUT/${TAG}_default.json; tail -3 $OUT/${TAG}_default.err
i=0
for kv in "$@"; do
  i=$((i+1))
  env $kv timeout 300 python tools/kernel_timing.py > $OUT/${TAG}_alt$i.json 2> $OUT/${TAG}_alt$i.err; cat $OUT/${TAG}_alt$i.json; tail -3 $OUT/${TAG}_alt$i.err
done
