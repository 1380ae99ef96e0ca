#!/bin/bash
# A/B: rounds kernel v3q vs v4q (incoherent launches), 8x4 pixel tiling on/off. Parity tests run on the new default.
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -x -q -m gpu > $OUT/ab2_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/ab2_pytest.log
show() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d["roofline"]
print(sys.argv[1], d["unit"], round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "ms/launch", round(r["ms_per_launch"],4), "nodes/ray", round(r["nodes_per_ray"],2), "share", d.get("extra",{}).get("trace_share_of_step"))
PY
}
for ik in v3q v4q v4; do
  TERRA_B200_INCOHERENT=$ik timeout 600 python bench.py --workload c3_1m --steps 3 --no-cpu-baseline > $OUT/ab2_c3_$ik.json 2>>$OUT/ab2_err.log; show $OUT/ab2_c3_$ik.json
done
for ik in v3q v4q; do for tile in 0 1; do
  TERRA_B200_INCOHERENT=$ik TERRA_B200_TILE=$tile timeout 600 python bench.py --steps 3 --no-cpu-baseline > $OUT/ab2_c2_${ik}_t$tile.json 2>>$OUT/ab2_err.log; show $OUT/ab2_c2_${ik}_t$tile.json
done; done
TERRA_B200_TRACE=v4 timeout 600 python bench.py --steps 3 --no-cpu-baseline > $OUT/ab2_c2_allv4.json 2>>$OUT/ab2_err.log; show $OUT/ab2_c2_allv4.json
tail -5 $OUT/ab2_err.log
