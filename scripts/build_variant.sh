#!/bin/sh
# A/B builds: scripts/build_variant.sh NAME "-DTB_...=..." -> terra_b200/ab/libterra_b200_NAME.so (loaded with
# TERRA_B200_LIB=...; git-ignored, travels to the GPU box). Same sources and flags as build.sh plus the defines.
set -e
name=$1; shift
cd "$(dirname "$0")/../terra_b200/csrc"
mkdir -p ../ab
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -fmad=false -std=c++17 \
     -Xcompiler -fPIC,-O2,-Wall -Xptxas -v "$@" -shared -o ../ab/libterra_b200_$name.so tb_render.cu tb_lbvh.cu tb_host.cpp tb_bvh.cpp \
     > ../ab/build_$name.log 2>&1
echo "built $name"
