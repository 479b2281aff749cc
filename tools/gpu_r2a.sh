#!/bin/bash
# Round-2 session A: parity of the new large-graph solver, graph timings per solver.
TAG=${1:-r02a}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_gpu.txt 2>&1; nproc >> $OUT/${TAG}_gpu.txt
echo "== pytest gn/dist"; timeout 1500 python -m pytest tests/test_gpu_gn.py tests/test_gpu_dist.py tests/test_gpu_global_opt.py -x -q > $OUT/${TAG}_pytest_gn.log 2>&1; echo "rc=$?"; tail -15 $OUT/${TAG}_pytest_gn.log
for S in auto dense pcg1; do
  echo "== graph k1024 solver=$S"
  B200SLAM_GN_SOLVER=$S timeout 900 python bench.py --workload graph --kfs 1024 --edges 8192 --steps 3 --warmup 2 > $OUT/${TAG}_graph_k1024_$S.json 2> $OUT/${TAG}_graph_k1024_$S.err; echo "rc=$?"
  python - <<PY
import json
try:
    d=json.load(open("$OUT/${TAG}_graph_k1024_$S.json"))["graph"]; print({k:d[k] for k in ("iters_per_s","ms_per_iter","phase_ms_per_iter","solver","data_gen_s")}, d["build_hbm"]["frac_of_peak"])
except Exception as e: print("ERR",e); print(open("$OUT/${TAG}_graph_k1024_$S.err").read()[-1500:])
PY
done
for CFG in "256 2048 rays" "64 512 rays" "32 128 calib" "32 128 rays"; do
  set -- $CFG
  for S in auto dense; do
    echo "== graph N=$1 E=$2 $3 solver=$S"
    B200SLAM_GN_SOLVER=$S timeout 600 python bench.py --workload graph --kfs $1 --edges $2 --kind $3 --steps 5 --warmup 3 > $OUT/${TAG}_graph_k$1_$3_$S.json 2> $OUT/${TAG}_graph_k$1_$3_$S.err; echo "rc=$?"
    python - <<PY
import json
try:
    d=json.load(open("$OUT/${TAG}_graph_k$1_$3_$S.json"))["graph"]; print({k:d[k] for k in ("iters_per_s","ms_per_iter","phase_ms_per_iter","solver")}, d["build_hbm"]["frac_of_peak"])
except Exception as e: print("ERR",e); print(open("$OUT/${TAG}_graph_k$1_$3_$S.err").read()[-1500:])
PY
  done
done
echo "== pytest rest"; timeout 1500 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_gn.py --deselect tests/test_gpu_dist.py --deselect tests/test_gpu_global_opt.py > $OUT/${TAG}_pytest_rest.log 2>&1; echo "rc=$?"; tail -5 $OUT/${TAG}_pytest_rest.log
