#!/bin/bash
# ncu --set full captures of the non-trace kernels of a C2 step: k_shade (both presets), k_raygen, k_accumulate.
TAG=${1:-r1w}; OUT=gpurun_out; mkdir -p $OUT
NCU_CMD="python bench.py --steps 1 --warmup 3 --spp 32 --no-cpu-baseline"
$NCU_CMD > $OUT/${TAG}_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:k_shade -s 10 -c 6 -f -o $OUT/${TAG}_k_shade $NCU_CMD > $OUT/${TAG}_ncu_shade.log 2>&1; echo "shade rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"k_raygen|k_accumulate" -s 2 -c 4 -f -o $OUT/${TAG}_k_raygen $NCU_CMD > $OUT/${TAG}_ncu_raygen.log 2>&1; echo "raygen rc=$?"
ls -la $OUT | grep $TAG
