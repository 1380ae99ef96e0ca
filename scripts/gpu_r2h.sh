#!/bin/bash
set -u
out=gpurun_out
python bench.py --impl reference --steps 1 --warmup 0 > $out/r2h_ref_c2.json 2> $out/r2h_ref_c2.err; echo "ref c2 rc=$?"; tail -c 700 $out/r2h_ref_c2.json
python bench.py --impl reference --workload c3_1m --steps 1 --warmup 0 > $out/r2h_ref_c3.json 2> $out/r2h_ref_c3.err; echo "ref c3 rc=$?"; tail -c 900 $out/r2h_ref_c3.json
python bench.py --workload c3_1m --steps 3 --warmup 3 > $out/r2h_bench_c3_1m.json 2> $out/r2h_bench_c3_1m.err; echo "c3_1m rc=$?"; tail -c 1200 $out/r2h_bench_c3_1m.json
python bench.py --workload c4 --spp 8 --steps 1 --warmup 3 --no-cpu-baseline > $out/r2h_bench_c4_small.json 2> $out/r2h_bench_c4_small.err; echo "c4 small rc=$?"; tail -c 600 $out/r2h_bench_c4_small.json; tail -3 $out/r2h_bench_c4_small.err
python bench.py --workload c1 --steps 5 --warmup 3 > $out/r2h_bench_c1.json 2> $out/r2h_bench_c1.err; echo "c1 rc=$?"; tail -c 400 $out/r2h_bench_c1.json
python bench.py --workload c5 --steps 2 --warmup 3 > $out/r2h_bench_c5.json 2> $out/r2h_bench_c5.err; echo "c5 rc=$?"; tail -c 400 $out/r2h_bench_c5.json
python bench.py --bvh reference --spp 4 --steps 2 --warmup 3 --no-cpu-baseline > $out/r2h_bench_c2_reftree.json 2> $out/r2h_bench_c2_reftree.err; echo "c2 reference tree rc=$?"; tail -c 400 $out/r2h_bench_c2_reftree.json
python bench.py --integrator 2 --steps 2 --warmup 3 --no-cpu-baseline > $out/r2h_bench_c2_mis.json 2> $out/r2h_bench_c2_mis.err; echo "c2 mis rc=$?"; tail -c 400 $out/r2h_bench_c2_mis.json
