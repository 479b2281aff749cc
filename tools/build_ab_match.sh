#!/bin/bash
# A/B build of the matching translation unit: bash tools/build_ab_match.sh <name> -DMACRO=... -> splatt3r_slam_b200/libb200slam_ab_<name>.so
set -e
NAME=$1; shift
cd "$(dirname "$0")/../splatt3r_slam_b200/csrc"
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC "$@" -c matching.cu -o /tmp/matching_ab_$NAME.o
nvcc -shared -o ../libb200slam_ab_$NAME.so /tmp/matching_ab_$NAME.o refine.o gn.o gs_raster.o 2>/dev/null
ls -la ../libb200slam_ab_$NAME.so
