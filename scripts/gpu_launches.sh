#!/bin/bash
# ncu launch list (per-launch gpu__time_duration) of the C2 bench at 32 spp (one wave per frame).
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-l1}
CMD="python bench.py --steps 1 --warmup 3 --spp 32 --no-cpu-baseline"
$CMD > $OUT/${TAG}_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu.log 2>&1
echo "rc=$?"; tail -1 $OUT/${TAG}_plain.log | cut -c1-200
