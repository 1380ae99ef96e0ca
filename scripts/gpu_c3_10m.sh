#!/bin/bash
# C3 at 10M triangles (HBM-resident geometry) on both builders; commit-stage timings at 1M and 10M.
OUT=gpurun_out; mkdir -p $OUT
for wl in c3_1m c3_10m; do for mode in sah3 lbvh; do
  TERRA_B200_BVH=$mode timeout 1500 python bench.py --workload $wl --steps 3 --no-cpu-baseline > $OUT/c3x_${wl}_$mode.json 2>>$OUT/c3x_err.log
  python - $OUT/c3x_${wl}_$mode.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d["roofline"]
print(sys.argv[1], round(d["value"],1), "Mrays/s frac", round(r["frac"],3), "nodes/ray", round(r["nodes_per_ray"],2), "commit", {k: round(v,1) for k,v in d["extra"]["commit_ms"].items()})
PY
done; done
tail -5 $OUT/c3x_err.log
