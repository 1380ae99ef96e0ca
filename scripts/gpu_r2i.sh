#!/bin/bash
# A/B of k_shade builds: default (prefetch, 3 CTAs/SM), no prefetch, 4 CTAs/SM (64 registers)
set -u
out=gpurun_out
for b in default nopf shade4; do
  if [ $b = default ]; then unset TERRA_B200_LIB; else export TERRA_B200_LIB=$PWD/terra_b200/libterra_b200_$b.so; fi
  python scripts/ab_trace.py c2 --only default --steps 5 >> $out/r2i_ab_shade.jsonl 2>> $out/r2i_ab_shade.err
  python scripts/ab_trace.py c5 --only default --spp 16 --steps 3 >> $out/r2i_ab_shade.jsonl 2>> $out/r2i_ab_shade.err
done
cat $out/r2i_ab_shade.jsonl
tail -3 $out/r2i_ab_shade.err
