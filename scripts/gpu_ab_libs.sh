#!/bin/bash
# A/B of library builds (scripts/build_variant.sh) on the C3 traversal benchmark and the C2 frame: every build under
# terra_b200/ab/ named on the command line, the variants of scripts/ab_trace.py whose name contains $ONLY, at the triangle
# counts in $TRIS. One JSON line per (build, variant, size).
set -u
TAG=$1; shift
OUT=gpurun_out; mkdir -p $OUT
TRIS=${TRIS:-"1000000 10000000"}
ONLY=${ONLY:-default}
for t in $TRIS; do
  for n in "$@"; do
    TERRA_B200_LIB=$PWD/terra_b200/ab/libterra_b200_$n.so timeout 600 python scripts/ab_trace.py c3 --tris $t --only "$ONLY" >> $OUT/${TAG}_ab_libs.jsonl 2>> $OUT/${TAG}_ab_libs.err
    echo "$n $t rc=$?"
  done
done
if [ -n "${C2:-}" ]; then
  for n in "$@"; do
    TERRA_B200_LIB=$PWD/terra_b200/ab/libterra_b200_$n.so timeout 600 python scripts/ab_trace.py c2 --steps 5 --only "default" >> $OUT/${TAG}_ab_libs.jsonl 2>> $OUT/${TAG}_ab_libs.err
    echo "$n c2 rc=$?"
  done
fi
if [ -n "${C5:-}" ]; then
  for n in "$@"; do
    TERRA_B200_LIB=$PWD/terra_b200/ab/libterra_b200_$n.so timeout 600 python scripts/ab_trace.py c5 --steps 5 --spp 16 --only "default" >> $OUT/${TAG}_ab_libs.jsonl 2>> $OUT/${TAG}_ab_libs.err
    echo "$n c5 rc=$?"
  done
fi
python - <<P
import json
for l in open("$OUT/${TAG}_ab_libs.jsonl"):
    d = json.loads(l)
    b = d['build'].split('libterra_b200_')[-1]
    if d['ab'] == 'c3':
        print(f"{b:20s} {d['variant'][:24]:24s} {d['tris']:>9d} {d['mrays_per_s']:8.1f} sum={d.get('hit_checksum')} nodes={d['nodes_per_ray']:.2f}")
    else:
        print(f"{b:20s} {d['ab']} {d['variant'][:24]:24s} {d['mpaths_per_s']:8.1f} Mpaths/s frame {d['ms']:.2f} ms trace {d['trace_ms_per_frame']:.2f} ms rest {d['ms'] - d['trace_ms_per_frame']:.2f} ms")
P
tail -3 $OUT/${TAG}_ab_libs.err
