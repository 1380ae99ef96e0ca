#!/bin/bash
# A/B of k_trace_v4 OPT bits (256-bit node loads, 10 CTAs/SM) and the rounds tunables on C3 1M; C2 on the best.
OUT=gpurun_out; mkdir -p $OUT
show() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d["roofline"]
print(sys.argv[1], d["unit"], round(d["value"],1), "ms/launch", round(r["ms_per_launch"],4), "nodes/ray", round(r["nodes_per_ray"],2), "share", d.get("extra",{}).get("trace_share_of_step"))
PY
}
for opt in 0 1 2 3; do
  TERRA_B200_V4_OPT=$opt timeout 600 python bench.py --workload c3_1m --steps 3 --no-cpu-baseline > $OUT/ab3_c3_o$opt.json 2>>$OUT/ab3_err.log; show $OUT/ab3_c3_o$opt.json
done
for tm in 4 12 16; do for rf in 22 26 30; do
  TERRA_B200_V4_OPT=${BEST_OPT:-1} TERRA_B200_TRI_MIN=$tm TERRA_B200_REFETCH=$rf timeout 600 python bench.py --workload c3_1m --steps 2 --rays 32000000 --no-cpu-baseline > $OUT/ab3_c3_t${tm}_r$rf.json 2>>$OUT/ab3_err.log; show $OUT/ab3_c3_t${tm}_r$rf.json
done; done
for opt in 0 1 3; do
  TERRA_B200_V4_OPT=$opt timeout 600 python bench.py --steps 3 --no-cpu-baseline > $OUT/ab3_c2_o$opt.json 2>>$OUT/ab3_err.log; show $OUT/ab3_c2_o$opt.json
done
TERRA_B200_V4_OPT=3 timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -2
tail -5 $OUT/ab3_err.log
