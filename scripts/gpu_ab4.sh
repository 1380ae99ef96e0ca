#!/bin/bash
# Packet kernel for camera rays: parity tests on the new default, then A/B against the per-ray kernel; v4 pop loop A/B.
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -x -q -m gpu > $OUT/ab4_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 $OUT/ab4_pytest.log
show() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d["roofline"]
print(sys.argv[1], d["unit"], round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "ms/launch", round(r["ms_per_launch"],4), "nodes/ray", round(r["nodes_per_ray"],2), "share", d.get("extra",{}).get("trace_share_of_step"))
PY
}
for ck in v2 packet; do
  TERRA_B200_COHERENT=$ck timeout 600 python bench.py --steps 3 --no-cpu-baseline > $OUT/ab4_c2_$ck.json 2>>$OUT/ab4_err.log; show $OUT/ab4_c2_$ck.json
done
for opt in 0 2 1 3; do
  TERRA_B200_V4_OPT=$opt timeout 600 python bench.py --workload c3_1m --steps 3 --no-cpu-baseline > $OUT/ab4_c3_o$opt.json 2>>$OUT/ab4_err.log; show $OUT/ab4_c3_o$opt.json
done
TERRA_B200_V4_OPT=2 timeout 600 python bench.py --steps 3 --no-cpu-baseline > $OUT/ab4_c2_packet_o2.json 2>>$OUT/ab4_err.log; show $OUT/ab4_c2_packet_o2.json
tail -5 $OUT/ab4_err.log
