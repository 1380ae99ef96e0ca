#!/bin/bash
# Sample-adjacent tiled slot layout + packet-kernel occupancy: parity tests, then C2 A/B.
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -x -q -m gpu > $OUT/ab5_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 $OUT/ab5_pytest.log
show() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d["roofline"]
print(sys.argv[1], d["unit"], round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "ms/launch", round(r["ms_per_launch"],4), "share", d.get("extra",{}).get("trace_share_of_step"))
PY
}
TERRA_B200_TILE=0 timeout 600 python bench.py --steps 3 --no-cpu-baseline > $OUT/ab5_c2_t0.json 2>>$OUT/ab5_err.log; show $OUT/ab5_c2_t0.json
for occ in 8 10 12; do
  TERRA_B200_PACKET_OCC=$occ timeout 600 python bench.py --steps 3 --no-cpu-baseline > $OUT/ab5_c2_occ$occ.json 2>>$OUT/ab5_err.log; show $OUT/ab5_c2_occ$occ.json
done
TERRA_B200_COHERENT=v2 timeout 600 python bench.py --steps 3 --no-cpu-baseline > $OUT/ab5_c2_v2.json 2>>$OUT/ab5_err.log; show $OUT/ab5_c2_v2.json
tail -5 $OUT/ab5_err.log
