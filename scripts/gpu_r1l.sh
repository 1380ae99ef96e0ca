#!/bin/bash
# r1l: graph replay on the device-resident path: GPU suite, smoke, C1 / C2 / C5 lines
OUT=gpurun_out; mkdir -p $OUT
timeout 1800 python -m pytest tests -m gpu -x -q > $OUT/r1l_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/r1l_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
for wl in c1 c2 c5; do
  timeout 900 python bench.py --workload $wl --no-cpu-baseline > $OUT/r1l_bench_$wl.json 2>>$OUT/r1l_err.log
  python - $OUT/r1l_bench_$wl.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], round(d["value"],1), d["unit"], "e2e", round(d["e2e"]["value"],1), "frac", round(d["roofline"]["frac"],3), "launches", d["gpu_launches"])
PY
done
tail -5 $OUT/r1l_err.log
