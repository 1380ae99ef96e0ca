#!/bin/bash
# ncu --set full of one kernel (regex $2) in the C2 bench at 32 spp.
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-n4}; K=${2:-k_trace_packet}; SKIP=${3:-3}
CMD="python bench.py --steps 1 --warmup 3 --spp 32 --no-cpu-baseline"
$CMD > $OUT/${TAG}_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$K -s $SKIP -c ${COUNT:-1} -f -o $OUT/${TAG}_$K $CMD > $OUT/${TAG}_ncu.log 2>&1
echo "rc=$?"
