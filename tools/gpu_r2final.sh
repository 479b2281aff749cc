#!/bin/bash
# Round-2 final session: full GPU tier, smoke, default bench line (all sub-blocks), reference arm, ncu launch list of the bench.
TAG=${1:-r02final}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_gpu.txt 2>&1; nproc >> $OUT/${TAG}_gpu.txt
echo "== pytest -m gpu"; timeout 600 python -m pytest tests -m gpu -q -x --durations=5 > $OUT/${TAG}_pytest.log 2>&1; echo "rc=$?"; tail -9 $OUT/${TAG}_pytest.log
echo "== smoke"; timeout 300 python __graft_entry__.py smoke > $OUT/${TAG}_smoke.log 2>&1; echo "rc=$?"; tail -1 $OUT/${TAG}_smoke.log
echo "== bench default"; ( time timeout 900 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err ) 2>&1 | grep real; tail -2 $OUT/${TAG}_bench.err
python - <<PY
import json
try:
    d=json.load(open("$OUT/${TAG}_bench.json"))
    for k in ("value","ms_per_step","e2e","kernel_us","roofline","graph","graph_cfg3","match_batched","ref_cuda","cpu_baseline","clocks","gpu_launches"):
        print(k, json.dumps(d.get(k))[:420])
except Exception as e: print("ERR",e)
PY
echo "== bench reference arm"; ( time timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench_reference.err ) 2>&1 | grep real; cut -c1-300 $OUT/${TAG}_bench_reference.json
echo "== ncu launch list"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file $OUT/${TAG}_launches.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_bench.log 2>&1
echo "launch list rc=$?"
