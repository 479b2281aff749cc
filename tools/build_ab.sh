#!/bin/bash
# A/B build of the refine kernel (tuning aid): bash tools/build_ab.sh <name> -DRT_TH=8 -DRT_BOX_KB=112 ...
# -> splatt3r_slam_b200/libb200slam_ab_<name>.so (select with B200SLAM_LIB=<path>); needs the default build's objects.
set -e
NAME=$1; shift
cd "$(dirname "$0")/../splatt3r_slam_b200/csrc"
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Xptxas -v -I. "$@" -c refine.cu -o /tmp/refine_ab_$NAME.o 2>&1 | grep -A2 "refine_tile" | grep -E "spill|Used"
nvcc -shared -o ../libb200slam_ab_$NAME.so matching.o /tmp/refine_ab_$NAME.o gn.o gs_raster.o 2>/dev/null
ls -la ../libb200slam_ab_$NAME.so
