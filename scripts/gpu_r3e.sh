#!/bin/bash
# Product build with k_trace_v8 (3 node steps per round, block refill) / k_trace_v7 (2 steps, block refill), refill below
# 24 lanes, triangle rounds from 6 parked leaves: whole GPU suite, A/B, bench lines, ncu of k_trace_v8 on C3 1M.
set -u
TAG=${1:-r3e}
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
C3="python bench.py --workload c3_1m --steps 1 --warmup 3 --rays 8000000 --no-cpu-baseline"
$C3 > $OUT/${TAG}_plain_c3.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_trace_v8 -s 3 -c 1 -f -o $OUT/${TAG}_c3_k_trace_v8 $C3 > $OUT/${TAG}_ncu_c3.log 2>&1
echo "ncu rc=$?"
ncu -i $OUT/${TAG}_c3_k_trace_v8.ncu-rep --page raw --csv > $OUT/${TAG}_c3_k_trace_v8_raw.csv 2> /dev/null
ncu -i $OUT/${TAG}_c3_k_trace_v8.ncu-rep --page source --csv 2> /dev/null | gzip > $OUT/${TAG}_c3_k_trace_v8_source.csv.gz
rm -f $OUT/${TAG}_*.ncu-rep
python scripts/round_stats.py > $OUT/${TAG}_round_stats.jsonl 2> $OUT/${TAG}_round_stats.err; echo "round stats rc=$?"
