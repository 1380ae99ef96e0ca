#!/bin/bash
# ncu --set full of the incoherent-ray kernel on the C3 1M workload (8M rays), source page imported.
OUT=gpurun_out; mkdir -p $OUT
TAG=${1:-n3}
C3="python bench.py --workload c3_1m --steps 1 --warmup 3 --rays 8000000 --no-cpu-baseline"
$C3 > $OUT/${TAG}_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_trace -s 3 -c 1 -f -o $OUT/${TAG}_c3 $C3 > $OUT/${TAG}_ncu.log 2>&1
echo "rc=$?"; tail -2 $OUT/${TAG}_plain.log | cut -c1-300
