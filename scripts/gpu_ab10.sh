#!/bin/bash
# A/B: per-preset register caps of k_shade (Diffuse / GGX at 3 or 4 CTAs per SM), same box.
OUT=gpurun_out; mkdir -p $OUT
for rep in 1 2; do for v in base light3 light4 all3; do
  if [ $v = base ]; then unset TERRA_B200_LIB; else export TERRA_B200_LIB=$PWD/terra_b200/libterra_b200_$v.so; fi
  for wl in c2 c5 c4; do
    [ $wl = c4 ] && [ $rep = 2 ] && continue
    timeout 900 python bench.py --workload $wl --steps 3 --no-cpu-baseline > $OUT/ab10_${wl}_${v}_$rep.json 2>>$OUT/ab10_err.log
    python - $OUT/ab10_${wl}_${v}_$rep.json $v $wl <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[3], sys.argv[2], round(d["value"],1), d["unit"])
PY
  done
done; done
tail -3 $OUT/ab10_err.log
