#!/bin/bash
# Round-2 closing captures at HEAD: the default bench line, the ncu launch list of a C2 step, --set full of the trace
# kernels (C2 wave, C3 1M, C3 10M) and of k_shade. Each ncu pass runs only after the same command exited 0 without ncu.
TAG=${1:-r2q}
OUT=gpurun_out; mkdir -p $OUT
python bench.py > $OUT/${TAG}_bench_c2.json 2> $OUT/${TAG}_bench_c2.err; echo "bench rc=$?"; tail -c 400 $OUT/${TAG}_bench_c2.json
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
ncu --set full --clock-control none --import-source on -k regex:k_trace_v6 -s 3 -c 1 -f -o $OUT/${TAG}_c3_k_trace $C3 > $OUT/${TAG}_ncu_c3.log 2>&1
echo "c3 rc=$?"
$C3B > $OUT/${TAG}_plain_c3_10m.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_trace_v6 -s 3 -c 1 -f -o $OUT/${TAG}_c3_10m_k_trace $C3B > $OUT/${TAG}_ncu_c3_10m.log 2>&1
echo "c3 10m rc=$?"
python bench.py --workload c3_10m --steps 3 --warmup 3 > $OUT/${TAG}_bench_c3_10m.json 2> $OUT/${TAG}_bench_c3_10m.err; echo "bench c3_10m rc=$?"; tail -5 $OUT/${TAG}_bench_c3_10m.err
# the reports are too large to travel (64 MiB limit): keep their raw pages as csv, and the source page of the C3 1M capture
for r in k_trace k_shade c3_k_trace c3_10m_k_trace; do
  ncu -i $OUT/${TAG}_$r.ncu-rep --page raw --csv > $OUT/${TAG}_${r}_raw.csv 2> /dev/null
done
ncu -i $OUT/${TAG}_c3_k_trace.ncu-rep --page source --csv 2> /dev/null | gzip > $OUT/${TAG}_c3_k_trace_source.csv.gz
rm -f $OUT/${TAG}_*.ncu-rep
ls -la $OUT | grep $TAG
