#!/bin/bash
set -u
out=gpurun_out
python scripts/ab_trace.py c3 > $out/r2e_ab_c3.jsonl 2> $out/r2e_ab_c3.err; echo "ab c3 rc=$?"
cut -c1-200 $out/r2e_ab_c3.jsonl
python scripts/ab_trace.py c3 --tris 10000000 --rays 16777216 > $out/r2e_ab_c3_10m.jsonl 2> $out/r2e_ab_c3_10m.err; echo "ab c3 10m rc=$?"
cut -c1-200 $out/r2e_ab_c3_10m.jsonl
python scripts/ab_trace.py c2 --spp 64 > $out/r2e_ab_c2.jsonl 2> $out/r2e_ab_c2.err; echo "ab c2 rc=$?"
cut -c1-330 $out/r2e_ab_c2.jsonl
