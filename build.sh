#!/bin/sh
# Builds terra_b200/libterra_b200.so (sm_100a only). -fmad=false: see tb_render.cu header.
set -e
cd "$(dirname "$0")/terra_b200/csrc"
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -fmad=false -std=c++17 \
     -Xcompiler -fPIC,-O2,-Wall ${TB_NVCC_EXTRA} -shared -o ../libterra_b200.so tb_render.cu tb_lbvh.cu tb_host.cpp tb_bvh.cpp
# headless front-end: the same C source against the product and, where the reference was
# compiled (oracle/_ref), against the unmodified reference — Terra API only.
cd ../..
mkdir -p tools/bin
${CC:-gcc} -std=gnu11 -O2 -Wall -DTERRA_HEADLESS_B200 -o tools/bin/terra_headless tools/terra_headless.c -Lterra_b200 -lterra_b200 -Wl,-rpath,'$ORIGIN/../../terra_b200' -lm
if [ -f oracle/_ref/libterra_ref.so ]; then
  ${CC:-gcc} -std=gnu11 -O2 -Wall -o tools/bin/terra_headless_ref tools/terra_headless.c oracle/_ref/libterra_ref.so -Wl,-rpath,'$ORIGIN/../../oracle/_ref' -lm
fi
