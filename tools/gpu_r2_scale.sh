#!/bin/bash
# Round-2 multi-GPU session: the driver's bench line at N = 2, 4, 8 ranks (tracker headline + configs[3] graph block).
TAG=${1:-r02s}; MAXN=${2:-8}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi topo -m > $OUT/${TAG}_topo.txt 2>&1
for N in 8 4 2; do
  [ $N -gt $MAXN ] && continue
  ( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --steps 20 --warmup 5 > $OUT/${TAG}_bench_n$N.json 2> $OUT/${TAG}_bench_n$N.err ) 2>&1 | grep real
  echo "N=$N rc=$?"; tail -n 2 $OUT/${TAG}_bench_n$N.err | cut -c1-300
  python - <<PY
import json
try:
    d=json.load(open("$OUT/${TAG}_bench_n$N.json")); g=d["graph"]
    print("N=$N value", d["value"], "ms/step", d["ms_per_step"], "e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e"].get("numa"))
    print("   graph", g["iters_per_s"], g["phase_ms_per_iter"], g["solver"]["cg_iters"], g["ranks_bit_identical"], g["pose_diff_vs_1gpu"], g["build_hbm"]["frac_of_peak"])
except Exception as e: print("ERR", e)
PY
done
