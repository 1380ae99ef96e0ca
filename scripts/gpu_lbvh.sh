#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_lbvh.py tests/test_gpu_parity.py -x -q -m gpu -s > $OUT/lbvh_pytest.log 2>&1; echo "pytest rc=$?"; grep -a "lbvh\|passed\|failed\|Error\|assert" $OUT/lbvh_pytest.log | tail -15
for mode in sah3 lbvh; do
  TERRA_B200_BVH=$mode timeout 900 python bench.py --workload c3_1m --steps 3 --no-cpu-baseline > $OUT/lbvh_c3_1m_$mode.json 2>>$OUT/lbvh_err.log
  python - $OUT/lbvh_c3_1m_$mode.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d["roofline"]
print(sys.argv[1], round(d["value"],1), "Mrays/s nodes/ray", round(r["nodes_per_ray"],2), "tris/ray", round(r["tris_per_ray"],2), "commit s", round(d["extra"]["scene_commit_s"],2))
PY
done
tail -5 $OUT/lbvh_err.log
