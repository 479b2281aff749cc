#!/bin/bash
# ncu --set full of the refine kernel (default library), raw csv exported on the box.
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-n}
timeout 300 python tools/prof_kernels.py > $OUT/${TAG}_plain_prof.log 2>&1 || { echo "plain run failed"; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'refine_tile_kernel|refine_qpq_kernel' -s 2 -c 1 \
    -o $OUT/${TAG}_refine -f python tools/prof_kernels.py > $OUT/${TAG}_ncu_refine.log 2>&1; echo "ncu rc=$?"
ncu -i $OUT/${TAG}_refine.ncu-rep --page raw --csv > $OUT/${TAG}_refine.csv 2>/dev/null
python tools/ncu_summary.py $OUT/${TAG}_refine.csv | grep -v "fp64\|_tc\|tmem\|tensor\|uniform\|lts__"
