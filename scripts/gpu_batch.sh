#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
for b in 4194304 16777216 67108864; do
  TERRA_B200_BATCH=$b timeout 600 python bench.py --steps 3 --no-cpu-baseline > $OUT/batch_$b.json 2>$OUT/batch_err.log
  python - <<PY
import json; d=json.load(open("$OUT/batch_$b.json")); print("batch $b", "Mpaths/s", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "trace ms/launch", round(d["roofline"]["ms_per_launch"],4), "rays/launch", int(d["roofline"]["rays_per_launch"]), "trace share", round(d["extra"]["trace_share_of_step"],3), "launches", d["gpu_launches"])
PY
done
tail -3 $OUT/batch_err.log
