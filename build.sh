#!/bin/sh
# Builds terra_b200/libterra_b200.so (sm_100a only). -fmad=false: see tb_render.cu header.
set -e
cd "$(dirname "$0")/terra_b200/csrc"
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -fmad=false -std=c++17 \
     -Xcompiler -fPIC,-O2,-Wall ${TB_NVCC_EXTRA} -shared -o ../libterra_b200.so tb_render.cu tb_lbvh.cu tb_host.cpp tb_bvh.cpp
