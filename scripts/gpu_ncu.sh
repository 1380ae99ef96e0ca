#!/bin/bash
# ncu --set full captures of k_trace: C2 frame (primary + bounce launches) and C3 incoherent rays.
TAG=${1:-n1}
OUT=gpurun_out; mkdir -p $OUT
C2="python bench.py --steps 1 --warmup 3 --spp 8 --no-cpu-baseline"
C3="python bench.py --workload c3_1m --steps 1 --warmup 3 --rays 8000000 --no-cpu-baseline"
$C2 > $OUT/${TAG}_plain_c2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_trace -s 20 -c 6 -f -o $OUT/${TAG}_c2_k_trace $C2 > $OUT/${TAG}_ncu_c2.log 2>&1
echo "c2 rc=$?"
$C3 > $OUT/${TAG}_plain_c3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_trace -s 3 -c 1 -f -o $OUT/${TAG}_c3_k_trace $C3 > $OUT/${TAG}_ncu_c3.log 2>&1
echo "c3 rc=$?"
ls -la $OUT | grep $TAG
