#!/bin/bash
# r1e: GPU tests, then C3 (1M, 10M) with the pipelined host-buffer path and the parallel host builder, chunk sweep.
OUT=gpurun_out; mkdir -p $OUT
nproc > $OUT/r1e_nproc.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/r1e_pytest.log 2>&1; echo "pytest exit $?" | tee -a $OUT/r1e_pytest.log
tail -3 $OUT/r1e_pytest.log
for wl in c3_1m c3_10m; do
  timeout 1500 python bench.py --workload $wl --steps 3 > $OUT/r1e_bench_$wl.json 2>>$OUT/r1e_err.log
  python - $OUT/r1e_bench_$wl.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d["roofline"]
print(sys.argv[1], round(d["value"],1), "Mrays/s e2e", d["e2e"], "commit", {k: round(v,1) for k,v in d["extra"]["commit_ms"].items()})
PY
done
for chunk in 262144 1048576 4194304 16777216; do
  TERRA_B200_IO_CHUNK=$chunk timeout 900 python bench.py --workload c3_1m --steps 2 --no-cpu-baseline > $OUT/r1e_chunk_$chunk.json 2>>$OUT/r1e_err.log
  python - $OUT/r1e_chunk_$chunk.json $chunk <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("chunk", sys.argv[2], "e2e", round(d["e2e"]["value"],1), "pageable", round(d["e2e"]["pageable_value"],1))
PY
done
timeout 900 python bench.py > $OUT/r1e_bench.json 2>>$OUT/r1e_err.log; tail -c 600 $OUT/r1e_bench.json
tail -5 $OUT/r1e_err.log
