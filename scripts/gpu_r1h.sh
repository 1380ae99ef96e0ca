#!/bin/bash
# r1h: merged counter kernels + staged pageable batches: GPU suite, tile pattern, C3 e2e (pinned / pageable), C2, C1.
OUT=gpurun_out; mkdir -p $OUT
timeout 1800 python -m pytest tests -m gpu -x -q > $OUT/r1h_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/r1h_pytest.log
for cfg in "128 8" "16 8"; do
  set -- $cfg
  for g in 0 default; do
    if [ $g = 0 ]; then export TERRA_B200_GRAPH_PATHS=0; else unset TERRA_B200_GRAPH_PATHS; fi
    timeout 600 python scripts/tile_bench.py $1 $2 3 2>>$OUT/r1h_err.log | tee -a $OUT/r1h_tiles.jsonl
  done
done
unset TERRA_B200_GRAPH_PATHS
for wl in c3_1m c1 c2 c5; do
  timeout 900 python bench.py --workload $wl > $OUT/r1h_bench_$wl.json 2>>$OUT/r1h_err.log
  python - $OUT/r1h_bench_$wl.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], round(d["value"],1), d["unit"], "e2e", {k: (round(v,1) if isinstance(v,float) else v) for k,v in d["e2e"].items() if k in ("value","pageable_value")}, "launches", d["gpu_launches"])
PY
done
tail -5 $OUT/r1h_err.log
