#!/bin/bash
# Round-2 session B: full GPU test tier, smoke, default bench line (all sub-blocks), reference arm.
TAG=${1:-r02b}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_gpu.txt 2>&1; nproc >> $OUT/${TAG}_gpu.txt; numactl -H >> $OUT/${TAG}_gpu.txt 2>&1
echo "== pytest -m gpu"; timeout 2400 python -m pytest tests -m gpu -q -x --durations=8 > $OUT/${TAG}_pytest.log 2>&1; echo "rc=$?"; tail -25 $OUT/${TAG}_pytest.log
echo "== smoke"; timeout 600 python __graft_entry__.py smoke > $OUT/${TAG}_smoke.log 2>&1; echo "rc=$?"; tail -2 $OUT/${TAG}_smoke.log
echo "== bench default"; ( time timeout 1200 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err ) 2>&1 | grep real; echo "rc=$?"; tail -3 $OUT/${TAG}_bench.err
python - <<PY
import json
try:
    d=json.load(open("$OUT/${TAG}_bench.json"))
    for k in ("value","ms_per_step","e2e","kernel_us","roofline","graph","graph_cfg3","match_batched","ref_cuda","cpu_baseline"):
        print(k, json.dumps(d.get(k))[:900])
except Exception as e: print("ERR",e)
PY
echo "== bench reference arm"; ( time timeout 900 python bench.py --impl reference --steps 5 --warmup 1 > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench_reference.err ) 2>&1 | grep real; cut -c1-300 $OUT/${TAG}_bench_reference.json
