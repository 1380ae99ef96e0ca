#!/bin/bash
# HEAD verification after k_trace_v7: whole GPU suite, smoke, the default bench line, the C3 lines, the v6 / v7 A/B at
# 1M / 3M / 10M triangles and one ncu --set full capture of k_trace_v7 on the 10M-triangle terrain.
set -u
TAG=${1:-r2u}
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -x -q -m gpu > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/${TAG}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/${TAG}_smoke.log
timeout 600 python bench.py > $OUT/${TAG}_bench_c2.json 2> $OUT/${TAG}_bench_c2.err; echo "bench rc=$?"; head -c 300 $OUT/${TAG}_bench_c2.json; echo
timeout 600 python bench.py --workload c3_1m --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_c3_1m.json 2> $OUT/${TAG}_bench_c3_1m.err; echo "c3_1m rc=$?"; head -c 200 $OUT/${TAG}_bench_c3_1m.json; echo
timeout 600 python bench.py --workload c3_10m --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_c3_10m.json 2> $OUT/${TAG}_bench_c3_10m.err; echo "c3_10m rc=$?"; head -c 200 $OUT/${TAG}_bench_c3_10m.json; echo
for t in 1000000 3000000 10000000; do
  timeout 600 python scripts/ab_trace.py c3 --tris $t --only "tree" >> $OUT/${TAG}_ab_c3_tree.jsonl 2>> $OUT/${TAG}_ab_c3_tree.err; echo "ab c3 $t rc=$?"
done
cat $OUT/${TAG}_ab_c3_tree.jsonl
C3B="python bench.py --workload c3_10m --steps 1 --warmup 3 --rays 8000000 --no-cpu-baseline"
$C3B > $OUT/${TAG}_plain_c3_10m.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_trace_v7 -s 3 -c 1 -f -o $OUT/${TAG}_c3_10m_k_trace_v7 $C3B > $OUT/${TAG}_ncu_c3_10m.log 2>&1
echo "ncu c3 10m rc=$?"
ncu -i $OUT/${TAG}_c3_10m_k_trace_v7.ncu-rep --page raw --csv > $OUT/${TAG}_c3_10m_k_trace_v7_raw.csv 2> /dev/null
rm -f $OUT/${TAG}_*.ncu-rep
ls -la $OUT | grep $TAG
