#!/bin/bash
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_lbvh.py tests/test_gpu_state.py tests/test_zz_no_alloc.py -x -q -m gpu > $OUT/r3c_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/r3c_pytest.log
for t in 1000000 10000000; do
timeout 900 python scripts/ab_trace.py c3 --tris $t --only "grid" >> $OUT/r3c_grid.jsonl 2>> $OUT/r3c_grid.err; echo "grid $t rc=$?"
done
python - <<P
import json
for l in open("$OUT/r3c_grid.jsonl"):
    d = json.loads(l)
    print(f"{d['variant']:32s} {d['tris']:>9d} {d['mrays_per_s']:8.1f} sum={d.get('hit_checksum')}")
P
