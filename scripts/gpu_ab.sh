#!/bin/bash
# A/B of trace-kernel variants: parity tests on the default, then short bench lines per variant.
TAG=${1:-ab}
OUT=gpurun_out; mkdir -p $OUT
timeout 1500 python -m pytest tests -x -q -m gpu > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/${TAG}_pytest.log
for v in ${VARIANTS:-v2 v1}; do
  echo "== $v c2"; TERRA_B200_TRACE=$v timeout 600 python bench.py --steps 3 --no-cpu-baseline > $OUT/${TAG}_c2_$v.json 2>$OUT/${TAG}_err.log; python - <<PY
import json; d=json.load(open("$OUT/${TAG}_c2_$v.json")); print("$v", "Mpaths/s", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "trace ms/launch", round(d["roofline"]["ms_per_launch"],4), "achieved GB/s", round(d["roofline"]["achieved"],1), "nodes/ray", round(d["roofline"]["nodes_per_ray"],2), "tris/ray", round(d["roofline"]["tris_per_ray"],2), "trace share", round(d["extra"]["trace_share_of_step"],3))
PY
  echo "== $v c3_1m"; TERRA_B200_TRACE=$v timeout 600 python bench.py --workload c3_1m --steps 3 --no-cpu-baseline > $OUT/${TAG}_c3_$v.json 2>>$OUT/${TAG}_err.log; python - <<PY
import json; d=json.load(open("$OUT/${TAG}_c3_$v.json")); print("$v", "Mrays/s", round(d["value"],1), "achieved GB/s", round(d["roofline"]["achieved"],1), "nodes/ray", round(d["roofline"]["nodes_per_ray"],2))
PY
done
tail -5 $OUT/${TAG}_err.log
