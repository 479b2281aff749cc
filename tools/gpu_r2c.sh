#!/bin/bash
# Round-2 session C: the allocation-free entry alone in a fresh process, the full GPU tier, smoke, the solver probe on
# ring-like graphs, the default bench line.
TAG=${1:-r02c}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_gpu.txt 2>&1; nproc >> $OUT/${TAG}_gpu.txt
echo "== no-alloc test alone"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > $OUT/${TAG}_noalloc.log 2>&1; echo "rc=$?"; tail -5 $OUT/${TAG}_noalloc.log
echo "== pytest -m gpu"; ( time timeout 2400 python -m pytest tests -m gpu -q -x --durations=8 > $OUT/${TAG}_pytest.log 2>&1 ) 2>&1 | grep real; tail -25 $OUT/${TAG}_pytest.log
echo "== smoke"; timeout 600 python __graft_entry__.py smoke > $OUT/${TAG}_smoke.log 2>&1; echo "rc=$?"; tail -2 $OUT/${TAG}_smoke.log
echo "== sparse probe"; timeout 600 python tools/sparse_probe.py > $OUT/${TAG}_sparse_probe.txt 2>&1; echo "rc=$?"; cat $OUT/${TAG}_sparse_probe.txt | tail -20
if [ "$2" != "nobench" ]; then
echo "== bench default"; ( time timeout 1200 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err ) 2>&1 | grep real; tail -3 $OUT/${TAG}_bench.err
python - <<PY
import json
try:
    d=json.load(open("$OUT/${TAG}_bench.json"))
    for k in ("value","ms_per_step","e2e","kernel_us","roofline","graph","graph_cfg3","match_batched","ref_cuda","cpu_baseline"):
        print(k, json.dumps(d.get(k))[:700])
except Exception as e: print("ERR",e)
PY
fi
