#!/bin/bash
# build an alternative product library with extra -D flags for gn.cu only:
#   tools/build_alt.sh NAME -DM3R_POLL_NS=20 ...   ->  tools/_build/NAME.so   (use with M3R_B200_LIB)
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p tools/_build
objs=lightweight-mast3r-slam_b200/build/libm3r_b200.so.o
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC "$@" -c -o tools/_build/$name.gn.o lightweight-mast3r-slam_b200/csrc/gn.cu
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o tools/_build/$name.so tools/_build/$name.gn.o $objs/common.cu.o $objs/matching.cu.o $objs/tracking.cu.o $objs/pointmap.cu.o
echo built tools/_build/$name.so
