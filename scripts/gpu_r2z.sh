#!/bin/bash
# k_trace_v8 / lean k_trace_v7 in the product build: whole GPU suite, A/B of the incoherent kernels on C3 (1M, 10M) and
# on the C2 / C5 frames, then the bench lines.
set -u
TAG=${1:-r2z}
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -x -q -m gpu > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/${TAG}_pytest_gpu.log
for t in 1000000 10000000; do
  timeout 600 python scripts/ab_trace.py c3 --tris $t --only "v" >> $OUT/${TAG}_ab_c3.jsonl 2>> $OUT/${TAG}_ab_c3.err; echo "ab c3 $t rc=$?"
done
timeout 600 python scripts/ab_trace.py c2 --steps 5 > $OUT/${TAG}_ab_c2.jsonl 2> $OUT/${TAG}_ab_c2.err; echo "ab c2 rc=$?"
timeout 600 python scripts/ab_trace.py c5 --steps 5 --spp 16 > $OUT/${TAG}_ab_c5.jsonl 2> $OUT/${TAG}_ab_c5.err; echo "ab c5 rc=$?"
python - <<P
import json
for f in ["c3", "c2", "c5"]:
    for l in open("$OUT/${TAG}_ab_%s.jsonl" % f):
        d = json.loads(l)
        if f == "c3":
            print(f"{d['variant'][:40]:40s} {d['tris']:>9d} {d['mrays_per_s']:8.1f} sum={d.get('hit_checksum')} same={d['hit_t_same_as_first_variant']} nodes={d['nodes_per_ray']:.2f}")
        else:
            print(f"{f} {d['variant'][:40]:40s} {d['mpaths_per_s']:8.1f} Mpaths/s  trace {d['trace_ms_per_frame']:.2f} ms  same={d['image_same_as_first_variant']}")
P
timeout 600 python bench.py > $OUT/${TAG}_bench_c2.json 2> $OUT/${TAG}_bench_c2.err; echo "bench rc=$?"; head -c 150 $OUT/${TAG}_bench_c2.json; echo
timeout 600 python bench.py --workload c3_1m --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_c3_1m.json 2> $OUT/${TAG}_bench_c3_1m.err; echo "c3_1m rc=$?"; head -c 150 $OUT/${TAG}_bench_c3_1m.json; echo
timeout 600 python bench.py --workload c3_10m --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_c3_10m.json 2> $OUT/${TAG}_bench_c3_10m.err; echo "c3_10m rc=$?"; head -c 150 $OUT/${TAG}_bench_c3_10m.json; echo
