#!/bin/bash
# r1f: new GPU tests (distribution, async jobs, env importance sampling, pipelined batches), then C2/C5 A/B of
# the shade kernel against the previous build in the same box.
OUT=gpurun_out; mkdir -p $OUT
timeout 1800 python -m pytest tests -m gpu -x -q -s > $OUT/r1f_pytest.log 2>&1; echo "pytest exit $?" | tee -a $OUT/r1f_pytest.log
grep -E "passed|failed|error|mean rel err|run-to-run" $OUT/r1f_pytest.log | tail -30
for rep in 1 2; do for lib in prev new; do
  if [ $lib = prev ]; then export TERRA_B200_LIB=$PWD/terra_b200/libterra_b200_prev.so; else unset TERRA_B200_LIB; fi
  for wl in c2 c5; do
    timeout 600 python bench.py --workload $wl --no-cpu-baseline > $OUT/r1f_${wl}_${lib}_$rep.json 2>>$OUT/r1f_err.log
    python - $OUT/r1f_${wl}_${lib}_$rep.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], round(d["value"],1), d["unit"], "e2e", round(d["e2e"]["value"],1))
PY
  done
done; done
tail -5 $OUT/r1f_err.log
