#!/bin/bash
# Shade-kernel register cap A/B (libs built with -DTB_SHADE_MIN_BLOCKS=3|4) on C2 and C5.
OUT=gpurun_out; mkdir -p $OUT
show() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d["roofline"]
print(sys.argv[1], d["unit"], round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "share", d.get("extra",{}).get("trace_share_of_step"))
PY
}
for v in default mb3 mb4; do
  if [ $v = default ]; then unset TERRA_B200_LIB; else export TERRA_B200_LIB=$PWD/gpurun_out_lib_$v.so; fi
  timeout 600 python bench.py --steps 3 --no-cpu-baseline > $OUT/ab8_c2_$v.json 2>>$OUT/ab8_err.log; show $OUT/ab8_c2_$v.json
  timeout 600 python bench.py --workload c5 --steps 2 --no-cpu-baseline > $OUT/ab8_c5_$v.json 2>>$OUT/ab8_err.log; show $OUT/ab8_c5_$v.json
done
tail -5 $OUT/ab8_err.log
