#!/bin/bash
# r1g: CUDA-graph replay of small terra_render calls: full GPU suite, then the satellite tile pattern with / without graphs.
OUT=gpurun_out; mkdir -p $OUT
timeout 1800 python -m pytest tests -m gpu -x -q > $OUT/r1g_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/r1g_pytest.log
for cfg in "128 8" "128 32" "64 8" "256 8" "16 8"; do
  set -- $cfg
  for g in 0 default; do
    if [ $g = 0 ]; then export TERRA_B200_GRAPH_PATHS=0; else unset TERRA_B200_GRAPH_PATHS; fi
    timeout 600 python scripts/tile_bench.py $1 $2 3 2>>$OUT/r1g_err.log | tee -a $OUT/r1g_tiles.jsonl
  done
done
unset TERRA_B200_GRAPH_PATHS
for wl in c1 c2; do
  timeout 600 python bench.py --workload $wl > $OUT/r1g_bench_$wl.json 2>>$OUT/r1g_err.log
  python - $OUT/r1g_bench_$wl.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], round(d["value"],1), d["unit"], "e2e", round(d["e2e"]["value"],1), "launches", d["gpu_launches"])
PY
done
tail -5 $OUT/r1g_err.log
