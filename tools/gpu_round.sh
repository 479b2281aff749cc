#!/bin/bash
# One GPU-box session: parity tests, bench lines, ncu launch list and full captures of the hot
# kernels.  Usage (from the repo root, under gpurun):  bash tools/gpu_round.sh <tag>
# Outputs land in gpurun_out/<tag>_*; copy the summaries you want kept into profiles/.
TAG=${1:-r01}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_gpu.txt 2>&1
nproc >> $OUT/${TAG}_gpu.txt
if [ "$2" != "full" ]; then
echo "== pytest -m gpu" ; timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/${TAG}_pytest.log
echo "== smoke" ; timeout 300 python __graft_entry__.py smoke > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/${TAG}_smoke.log
echo "== bench tracker"; timeout 600 python bench.py > $OUT/${TAG}_bench_tracker.json 2> $OUT/${TAG}_bench_tracker.err; echo "rc=$?"; cat $OUT/${TAG}_bench_tracker.json
echo "== bench graph";   timeout 600 python bench.py --workload graph --steps 10 > $OUT/${TAG}_bench_graph.json 2> $OUT/${TAG}_bench_graph.err; echo "rc=$?"; cat $OUT/${TAG}_bench_graph.json
echo "== bench graph N=256";  timeout 600 python bench.py --workload graph --kfs 256 --edges 2048 --steps 3 --warmup 3 > $OUT/${TAG}_bench_graph_k256.json 2> $OUT/${TAG}_bench_graph_k256.err; echo "rc=$?"; cut -c1-200 $OUT/${TAG}_bench_graph_k256.json
echo "== bench graph N=1024 (configs[3])"; timeout 900 python bench.py --workload graph --kfs 1024 --edges 8192 --steps 3 --warmup 3 > $OUT/${TAG}_bench_graph_k1024.json 2> $OUT/${TAG}_bench_graph_k1024.err; echo "rc=$?"; cut -c1-200 $OUT/${TAG}_bench_graph_k1024.json
echo "== kernel timing"; timeout 600 python tools/kernel_timing.py > $OUT/${TAG}_kernel_timing.json 2> $OUT/${TAG}_kernel_timing.err; echo "rc=$?"; tail -c 1500 $OUT/${TAG}_kernel_timing.json
echo "== bench reference"; timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench_reference.err; echo "rc=$?"; cat $OUT/${TAG}_bench_reference.json
echo "== parity report"; timeout 900 python tools/parity_report.py > $OUT/${TAG}_parity.json 2> $OUT/${TAG}_parity.err; echo "rc=$?"
fi
# at most ONE profiler invocation per box session: "list" = launch list of the bench, "full" = --set full capture
if [ "$2" == "list" ]; then
echo "== ncu launch list (bench, short)"
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_plain_bench.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/${TAG}_launches.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_bench.log 2>&1
echo "launch list rc=$?"
fi
if [ "$2" == "full" ]; then
echo "== ncu full (hot kernels)"
timeout 300 python tools/prof_kernels.py > $OUT/${TAG}_plain_prof.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on \
    -k regex:'iter_proj_kernel|refine_tile_kernel|gn_persistent_kernel|prep_planes_kernel|match_proj_kernel|track_kernel' -s 12 -c 8 \
    -o $OUT/${TAG}_prof -f python tools/prof_kernels.py > $OUT/${TAG}_ncu_prof.log 2>&1
echo "ncu full rc=$?"; tail -3 $OUT/${TAG}_ncu_prof.log
fi
ls -la $OUT
