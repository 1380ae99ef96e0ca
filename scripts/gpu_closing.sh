#!/bin/bash
# Closing captures at HEAD: whole GPU suite, smoke, the default bench line and the other workloads' lines, the ncu launch
# list of a C2 step, --set full of the trace kernels (C2 wave, C3 1M, C3 10M) and of k_shade. Each ncu pass runs only
# after the same command exited 0 without ncu. scripts/make_traffic.py + scripts/summarize_ncu.py turn the output into
# profiles/traffic.json and the profiles/<tag>_*_metrics.md summaries.
TAG=${1:-r3g}
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -x -q -m gpu > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/${TAG}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $OUT/${TAG}_smoke.log
timeout 600 python bench.py > $OUT/${TAG}_bench_c2.json 2> $OUT/${TAG}_bench_c2.err; echo "bench rc=$?"; head -c 120 $OUT/${TAG}_bench_c2.json; echo
timeout 600 python bench.py --impl reference > $OUT/${TAG}_ref_c2.json 2> $OUT/${TAG}_ref_c2.err; echo "ref rc=$?"; head -c 120 $OUT/${TAG}_ref_c2.json; echo
for wl in c1 c5; do
  timeout 600 python bench.py --workload $wl --no-cpu-baseline > $OUT/${TAG}_bench_$wl.json 2> $OUT/${TAG}_bench_$wl.err; echo "$wl rc=$?"; head -c 120 $OUT/${TAG}_bench_$wl.json; echo
done
timeout 900 python bench.py --workload c3_1m --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_c3_1m.json 2> $OUT/${TAG}_bench_c3_1m.err; echo "c3_1m rc=$?"; head -c 120 $OUT/${TAG}_bench_c3_1m.json; echo
timeout 900 python bench.py --workload c3_10m --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_c3_10m.json 2> $OUT/${TAG}_bench_c3_10m.err; echo "c3_10m rc=$?"; head -c 120 $OUT/${TAG}_bench_c3_10m.json; echo
C2="python bench.py --steps 1 --warmup 3 --spp 32 --no-cpu-baseline"
C3="python bench.py --workload c3_1m --steps 1 --warmup 3 --rays 8000000 --no-cpu-baseline"
C3B="python bench.py --workload c3_10m --steps 1 --warmup 3 --rays 8000000 --no-cpu-baseline"
$C2 > $OUT/${TAG}_plain_c2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 140 -c 400 --csv --log-file $OUT/${TAG}_launches.csv $C2 > $OUT/${TAG}_ncu_launches.log 2>&1
echo "launches rc=$?"
$C2 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_trace -s 20 -c 5 -f -o $OUT/${TAG}_k_trace $C2 > $OUT/${TAG}_ncu_c2.log 2>&1
echo "c2 trace rc=$?"
$C2 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_shade -s 16 -c 4 -f -o $OUT/${TAG}_k_shade $C2 > $OUT/${TAG}_ncu_shade.log 2>&1
echo "c2 shade rc=$?"
$C3 > $OUT/${TAG}_plain_c3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_trace_v8 -s 3 -c 1 -f -o $OUT/${TAG}_c3_k_trace $C3 > $OUT/${TAG}_ncu_c3.log 2>&1
echo "c3 rc=$?"
$C3B > $OUT/${TAG}_plain_c3_10m.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_trace_v7 -s 3 -c 1 -f -o $OUT/${TAG}_c3_10m_k_trace $C3B > $OUT/${TAG}_ncu_c3_10m.log 2>&1
echo "c3 10m rc=$?"
# the reports are too large to travel (64 MiB limit): keep their raw pages as csv, and the source page of the C3 1M capture
for r in k_trace k_shade c3_k_trace c3_10m_k_trace; do
  ncu -i $OUT/${TAG}_$r.ncu-rep --page raw --csv > $OUT/${TAG}_${r}_raw.csv 2> /dev/null
done
ncu -i $OUT/${TAG}_c3_k_trace.ncu-rep --page source --csv 2> /dev/null | gzip > $OUT/${TAG}_c3_k_trace_source.csv.gz
rm -f $OUT/${TAG}_*.ncu-rep
python scripts/round_stats.py > $OUT/${TAG}_round_stats.jsonl 2> $OUT/${TAG}_round_stats.err; echo "round stats rc=$?"
ls -la $OUT | grep $TAG | wc -l
