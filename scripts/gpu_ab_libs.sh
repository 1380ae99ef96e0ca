#!/bin/bash
# A/B of library builds (scripts/build_variant.sh) on the C3 traversal benchmark: every build under terra_b200/ab/ named
# on the command line, v7 and v6 each, at the triangle counts in $TRIS. One JSON line per (build, kernel, size).
set -u
TAG=$1; shift
OUT=gpurun_out; mkdir -p $OUT
TRIS=${TRIS:-"1000000 10000000"}
ONLY=${ONLY:-tree}
for t in $TRIS; do
  for n in "$@"; do
    TERRA_B200_LIB=$PWD/terra_b200/ab/libterra_b200_$n.so timeout 600 python scripts/ab_trace.py c3 --tris $t --only "$ONLY" >> $OUT/${TAG}_ab_libs.jsonl 2>> $OUT/${TAG}_ab_libs.err
    echo "$n $t rc=$?"
  done
done
python - <<P
import json
for l in open("$OUT/${TAG}_ab_libs.jsonl"):
    d = json.loads(l)
    print(f"{d['build'].split('libterra_b200_')[-1]:24s} {d['variant'][:20]:20s} {d['tris']:>9d} {d['mrays_per_s']:8.1f} same={d['hit_t_same_as_first_variant']} sum={d.get('hit_checksum')} nodes={d['nodes_per_ray']:.2f}")
P
tail -3 $OUT/${TAG}_ab_libs.err
