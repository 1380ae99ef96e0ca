#!/bin/bash
OUT=gpurun_out
for tm in 4 8 16 24; do for rf in 16 20 26 30; do
  r=$(TERRA_B200_TRI_MIN=$tm TERRA_B200_REFETCH=$rf python bench.py --workload c3_1m --steps 2 --rays 32000000 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['roofline']['nodes_per_ray'],2))")
  echo "tri_min $tm refetch $rf : $r"
done; done
