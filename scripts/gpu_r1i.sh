#!/bin/bash
# r1i: where a small terra_render call spends its time (bounces sweep), staged pageable batches with a touched output
OUT=gpurun_out; mkdir -p $OUT
for b in 0 1 2 4; do
  for g in 0 default; do
    if [ $g = 0 ]; then export TERRA_B200_GRAPH_PATHS=0; else unset TERRA_B200_GRAPH_PATHS; fi
    timeout 600 python scripts/tile_bench.py 128 8 3 $b 2>>$OUT/r1i_err.log | tee -a $OUT/r1i_tiles.jsonl
  done
done
unset TERRA_B200_GRAPH_PATHS
for sm in default 0; do
  if [ $sm = 0 ]; then export TERRA_B200_STAGING_MIN=0; else unset TERRA_B200_STAGING_MIN; fi
  timeout 900 python bench.py --workload c3_1m --steps 2 --no-cpu-baseline > $OUT/r1i_c3_staging_$sm.json 2>>$OUT/r1i_err.log
  python - $OUT/r1i_c3_staging_$sm.json $sm <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("staging_min", sys.argv[2], "e2e pinned", round(d["e2e"]["value"],1), "pageable", round(d["e2e"]["pageable_value"],1))
PY
done
tail -5 $OUT/r1i_err.log
