#!/bin/bash
# Samples-per-warp wave layout A/B (TERRA_B200_WARP_SAMPLES = 1 | 4 | 8 | 32) on C2, C5 and C4-like 4K; parity suite first.
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -x -q -m gpu > $OUT/ab6_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/ab6_pytest.log
show() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d["roofline"]
print(sys.argv[1], d["unit"], round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "ms/launch", round(r["ms_per_launch"],4), "share", d.get("extra",{}).get("trace_share_of_step"))
PY
}
for ws in 1 4 8 32; do
  TERRA_B200_WARP_SAMPLES=$ws timeout 600 python bench.py --steps 3 --no-cpu-baseline > $OUT/ab6_c2_ws$ws.json 2>>$OUT/ab6_err.log; show $OUT/ab6_c2_ws$ws.json
done
for ws in 1 32; do
  TERRA_B200_WARP_SAMPLES=$ws timeout 600 python bench.py --workload c5 --steps 2 --no-cpu-baseline > $OUT/ab6_c5_ws$ws.json 2>>$OUT/ab6_err.log; show $OUT/ab6_c5_ws$ws.json
done
tail -5 $OUT/ab6_err.log
