#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
C3="python bench.py --workload c3_1m --steps 1 --warmup 3 --rays 8000000 --no-cpu-baseline"
for v in ${VARIANTS:-v3q}; do
  TERRA_B200_TRACE=$v $C3 > $OUT/n2_plain_$v.log 2>&1 && \
  TERRA_B200_TRACE=$v ncu --set full --clock-control none --import-source on -k regex:k_trace -s 3 -c 1 -f -o $OUT/n2_c3_$v $C3 > $OUT/n2_ncu_$v.log 2>&1
  echo "$v rc=$?"
done
