#!/bin/bash
# Round 2, first GPU call: parity suite + smoke + the contract bench line + A/B of the trace kernels.
set -u
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/r2a_gpu.txt 2>&1
nproc >> $out/r2a_gpu.txt
python -m pytest tests -m gpu -x -q -s > $out/r2a_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $out/r2a_gpu.txt
tail -5 $out/r2a_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $out/r2a_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $out/r2a_gpu.txt
tail -3 $out/r2a_smoke.log
python bench.py --steps 3 --warmup 3 > $out/r2a_bench.json 2> $out/r2a_bench.err; echo "bench rc=$?" | tee -a $out/r2a_gpu.txt
tail -c 1500 $out/r2a_bench.json
python scripts/ab_trace.py c3 > $out/r2a_ab_c3.jsonl 2> $out/r2a_ab_c3.err; echo "ab c3 rc=$?" | tee -a $out/r2a_gpu.txt
cat $out/r2a_ab_c3.jsonl
TERRA_B200_LIB=$PWD/terra_b200/libterra_b200_occ7.so python scripts/ab_trace.py c3 --only "v6 opt" > $out/r2a_ab_c3_occ7.jsonl 2> $out/r2a_ab_c3_occ7.err; echo "ab c3 occ7 rc=$?" | tee -a $out/r2a_gpu.txt
cat $out/r2a_ab_c3_occ7.jsonl
python scripts/ab_trace.py c2 --spp 64 > $out/r2a_ab_c2.jsonl 2> $out/r2a_ab_c2.err; echo "ab c2 rc=$?" | tee -a $out/r2a_gpu.txt
cat $out/r2a_ab_c2.jsonl
TERRA_B200_LIB=$PWD/terra_b200/libterra_b200_occ7.so python scripts/ab_trace.py c2 --spp 64 --only "v6 opt1 + fused raygen" > $out/r2a_ab_c2_occ7.jsonl 2> $out/r2a_ab_c2_occ7.err
cat $out/r2a_ab_c2_occ7.jsonl
