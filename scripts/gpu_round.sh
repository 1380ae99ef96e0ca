#!/bin/bash
# One GPU-box session: parity tests, smoke, the bench contract, then the two ncu passes
# (launch list + one --set full capture of the dominant kernel). Outputs land in gpurun_out/.
# usage (from the repo root, under gpurun): bash scripts/gpu_round.sh [tag]
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi -L > $OUT/${TAG}_gpu.txt; nproc >> $OUT/${TAG}_gpu.txt
echo "== pytest -m gpu"; timeout 1500 python -m pytest tests -x -q -m gpu > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $OUT/${TAG}_pytest_gpu.log
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 $OUT/${TAG}_smoke.log
echo "== bench"; timeout 900 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"; cat $OUT/${TAG}_bench.json
echo "== bench reference arm"; timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/${TAG}_bench_ref.json 2>> $OUT/${TAG}_bench.err; cat $OUT/${TAG}_bench_ref.json
echo "== bench c3_1m"; timeout 900 python bench.py --workload c3_1m --steps 3 > $OUT/${TAG}_bench_c3_1m.json 2>> $OUT/${TAG}_bench.err; cat $OUT/${TAG}_bench_c3_1m.json
echo "== bench c3_10m"; timeout 1500 python bench.py --workload c3_10m --steps 3 --no-cpu-baseline > $OUT/${TAG}_bench_c3_10m.json 2>> $OUT/${TAG}_bench.err; cat $OUT/${TAG}_bench_c3_10m.json
echo "== bench c4"; timeout 1500 python bench.py --workload c4 --steps 2 --no-cpu-baseline > $OUT/${TAG}_bench_c4.json 2>> $OUT/${TAG}_bench.err; cat $OUT/${TAG}_bench_c4.json
echo "== bench c5 / c1"; timeout 900 python bench.py --workload c5 --steps 2 --no-cpu-baseline > $OUT/${TAG}_bench_c5.json 2>> $OUT/${TAG}_bench.err; cat $OUT/${TAG}_bench_c5.json
timeout 600 python bench.py --workload c1 --steps 5 --no-cpu-baseline > $OUT/${TAG}_bench_c1.json 2>> $OUT/${TAG}_bench.err; cat $OUT/${TAG}_bench_c1.json
NCU_CMD="python bench.py --steps 1 --warmup 3 --spp 32 --no-cpu-baseline"
echo "== ncu launch list"
$NCU_CMD > $OUT/${TAG}_ncu_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/${TAG}_launches.csv $NCU_CMD > $OUT/${TAG}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
echo "== ncu full capture of k_trace"
ncu --set full --clock-control none --import-source on -k regex:k_trace -s 10 -c 5 -f -o $OUT/${TAG}_k_trace $NCU_CMD > $OUT/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la $OUT | head -40
