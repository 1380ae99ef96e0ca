#!/bin/bash
# A/B: direction binning of the bounce queues (TERRA_B200_SORT), same box. Parity first.
OUT=gpurun_out; mkdir -p $OUT
TERRA_B200_SORT=2 timeout 900 python -m pytest tests/test_gpu_render.py tests/test_gpu_parity.py -m gpu -x -q > $OUT/ab9_pytest.log 2>&1; echo "pytest(sort=2) exit $?"; tail -2 $OUT/ab9_pytest.log
run() { # name, env...
  name=$1; shift
  for wl in c2 c5; do
    env "$@" timeout 600 python bench.py --workload $wl --no-cpu-baseline > $OUT/ab9_${wl}_$name.json 2>>$OUT/ab9_err.log
    python - $OUT/ab9_${wl}_$name.json $name $wl <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[3], sys.argv[2], round(d["value"],1), d["unit"], "launches", d["gpu_launches"])
PY
  done
}
run off TERRA_B200_SORT=0
run oct_b1 TERRA_B200_SORT=1 TERRA_B200_SORT_BOUNCES=1
run oct_b2 TERRA_B200_SORT=1 TERRA_B200_SORT_BOUNCES=2
run oct_b9 TERRA_B200_SORT=1 TERRA_B200_SORT_BOUNCES=99
run octax_b1 TERRA_B200_SORT=2 TERRA_B200_SORT_BOUNCES=1
run octax_b2 TERRA_B200_SORT=2 TERRA_B200_SORT_BOUNCES=2
run off2 TERRA_B200_SORT=0
tail -3 $OUT/ab9_err.log
