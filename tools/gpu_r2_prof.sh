#!/bin/bash
# Round-2 profiling session (one GPU): launch list of the bench + ncu --set full captures of the shipped kernels.
TAG=${1:-r02p}; OUT=gpurun_out; mkdir -p $OUT
NCU="ncu --set full --clock-control none --import-source on"
echo "== launch list"
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_plain_bench.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $OUT/${TAG}_launches.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_bench.log 2>&1; echo "rc=$?"
echo "== tracker kernels"
timeout 300 python tools/prof_kernels.py > $OUT/${TAG}_plain_prof.log 2>&1 &&
timeout 1500 $NCU -k regex:'iter_proj_kernel|refine_tile_kernel|gn_persistent_kernel|prep_planes_kernel|match_proj_kernel|track_kernel' -s 12 -c 8 \
    -o $OUT/${TAG}_tracker -f python tools/prof_kernels.py > $OUT/${TAG}_ncu_tracker.log 2>&1; echo "rc=$?"
echo "== graph32"
timeout 300 python tools/prof_r2.py graph32 > $OUT/${TAG}_plain_g32.log 2>&1 &&
timeout 1500 $NCU -k regex:'build_kernel|solve_retract_kernel|finalize_kernel|assemble_kernel' -s 8 -c 8 \
    -o $OUT/${TAG}_graph32 -f python tools/prof_r2.py graph32 > $OUT/${TAG}_ncu_g32.log 2>&1; echo "rc=$?"
echo "== graph256"
timeout 300 python tools/prof_r2.py graph256 > $OUT/${TAG}_plain_g256.log 2>&1 &&
timeout 1500 $NCU -k regex:'build_kernel|pcg2_solve_retract_kernel' -s 4 -c 4 \
    -o $OUT/${TAG}_graph256 -f python tools/prof_r2.py graph256 > $OUT/${TAG}_ncu_g256.log 2>&1; echo "rc=$?"
echo "== solver1024"
timeout 300 python tools/prof_r2.py solver1024 > $OUT/${TAG}_plain_s1024.log 2>&1 &&
timeout 1500 $NCU -k regex:'pcg2_solve_retract_kernel' -s 1 -c 1 \
    -o $OUT/${TAG}_solver1024 -f python tools/prof_r2.py solver1024 > $OUT/${TAG}_ncu_s1024.log 2>&1; echo "rc=$?"
echo "== tcgen05 dense-correlation experiment"
timeout 120 ./tools/micro/tc_corr > $OUT/${TAG}_tc_corr.txt 2>&1 &&
timeout 600 $NCU -k regex:'corr_kernel' -s 2 -c 3 -o $OUT/${TAG}_tccorr -f ./tools/micro/tc_corr > $OUT/${TAG}_ncu_tccorr.log 2>&1; echo "rc=$?"
# the reports are large (gpurun_out merges at most 64 MiB): export the raw counter pages here and drop the reports
for R in tracker graph32 graph256 solver1024 tccorr; do
  [ -f $OUT/${TAG}_$R.ncu-rep ] && ncu -i $OUT/${TAG}_$R.ncu-rep --page raw --csv > $OUT/${TAG}_${R}_raw.csv 2>/dev/null
  [ "$KEEP_REP" == "$R" ] || rm -f $OUT/${TAG}_$R.ncu-rep
done
ls -la $OUT | grep $TAG
