#!/bin/bash
set -u
out=gpurun_out
python scripts/ab_trace.py c3 > $out/r2f_ab_c3.jsonl 2> $out/r2f_ab_c3.err; echo "ab c3 rc=$?"
cut -c1-200 $out/r2f_ab_c3.jsonl
TERRA_B200_LIB=$PWD/terra_b200/libterra_b200_occ8.so python scripts/ab_trace.py c3 --only "opt" > $out/r2f_ab_c3_occ8.jsonl 2> $out/r2f_ab_c3_occ8.err
cut -c1-220 $out/r2f_ab_c3_occ8.jsonl
python scripts/ab_trace.py c3 --tris 10000000 --rays 16777216 > $out/r2f_ab_c3_10m.jsonl 2> $out/r2f_ab_c3_10m.err; echo "ab c3 10m rc=$?"
cut -c1-200 $out/r2f_ab_c3_10m.jsonl
python scripts/ab_trace.py c2 --spp 64 > $out/r2f_ab_c2.jsonl 2> $out/r2f_ab_c2.err; echo "ab c2 rc=$?"
cut -c1-330 $out/r2f_ab_c2.jsonl
TERRA_B200_LIB=$PWD/terra_b200/libterra_b200_occ8.so python scripts/ab_trace.py c2 --spp 64 --only "opt16" > $out/r2f_ab_c2_occ8.jsonl 2>&1
cut -c1-330 $out/r2f_ab_c2_occ8.jsonl
