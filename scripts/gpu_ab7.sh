#!/bin/bash
# 4-wide tree kernel (k_trace_v5) vs the binary rounds kernel: parity suite on v5, then A/B.
OUT=gpurun_out; mkdir -p $OUT
TERRA_B200_INCOHERENT=v5 timeout 900 python -m pytest tests -x -q -m gpu > $OUT/ab7_pytest.log 2>&1; echo "pytest(v5) rc=$?"; tail -6 $OUT/ab7_pytest.log
show() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d["roofline"]
print(sys.argv[1], d["unit"], round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "ms/launch", round(r["ms_per_launch"],4), "share", d.get("extra",{}).get("trace_share_of_step"), "agree", d.get("extra",{}).get("hit_t_agreement_with_oracle"))
PY
}
for ik in v4q v5; do
  TERRA_B200_INCOHERENT=$ik timeout 600 python bench.py --workload c3_1m --steps 3 > $OUT/ab7_c3_$ik.json 2>>$OUT/ab7_err.log; show $OUT/ab7_c3_$ik.json
  TERRA_B200_INCOHERENT=$ik timeout 600 python bench.py --steps 3 --no-cpu-baseline > $OUT/ab7_c2_$ik.json 2>>$OUT/ab7_err.log; show $OUT/ab7_c2_$ik.json
done
for tm in 4 12; do
  TERRA_B200_INCOHERENT=v5 TERRA_B200_TRI_MIN=$tm timeout 600 python bench.py --workload c3_1m --steps 3 --no-cpu-baseline > $OUT/ab7_c3_v5_t$tm.json 2>>$OUT/ab7_err.log; show $OUT/ab7_c3_v5_t$tm.json
done
tail -5 $OUT/ab7_err.log
