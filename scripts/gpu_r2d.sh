#!/bin/bash
# ncu captures at HEAD: launch list of a C2 step, --set full of the trace kernels (C2 wave, C3 1M) and of k_shade.
TAG=${1:-r2d}
OUT=gpurun_out; mkdir -p $OUT
C2="python bench.py --steps 1 --warmup 3 --spp 32 --no-cpu-baseline"
C3="python bench.py --workload c3_1m --steps 1 --warmup 3 --rays 8000000 --no-cpu-baseline"
$C2 > $OUT/${TAG}_plain_c2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 140 -c 400 --csv --log-file $OUT/${TAG}_launches.csv $C2 > $OUT/${TAG}_ncu_launches.log 2>&1
echo "launches rc=$?"
$C2 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_trace -s 20 -c 5 -f -o $OUT/${TAG}_k_trace $C2 > $OUT/${TAG}_ncu_c2.log 2>&1
echo "c2 trace rc=$?"
$C2 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_shade -s 16 -c 4 -f -o $OUT/${TAG}_k_shade $C2 > $OUT/${TAG}_ncu_shade.log 2>&1
echo "c2 shade rc=$?"
$C3 > $OUT/${TAG}_plain_c3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_trace_v6 -s 3 -c 1 -f -o $OUT/${TAG}_c3_k_trace $C3 > $OUT/${TAG}_ncu_c3.log 2>&1
echo "c3 rc=$?"
ls -la $OUT | grep $TAG
tail -c 600 $OUT/${TAG}_plain_c3.log
