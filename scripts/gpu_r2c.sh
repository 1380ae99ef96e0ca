#!/bin/bash
set -u
out=gpurun_out
python -m pytest tests -m gpu -q > $out/r2c_pytest_gpu.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|FAILED|ERROR" $out/r2c_pytest_gpu.log | tail -10
python scripts/ab_trace.py c3 > $out/r2c_ab_c3.jsonl 2> $out/r2c_ab_c3.err; echo "ab c3 rc=$?"
cut -c1-200 $out/r2c_ab_c3.jsonl
TERRA_B200_LIB=$PWD/terra_b200/libterra_b200_occ8.so python scripts/ab_trace.py c3 --only "v6 opt" > $out/r2c_ab_c3_occ8.jsonl 2> $out/r2c_ab_c3_occ8.err
cut -c1-200 $out/r2c_ab_c3_occ8.jsonl
python scripts/ab_trace.py c2 --spp 64 > $out/r2c_ab_c2.jsonl 2> $out/r2c_ab_c2.err; echo "ab c2 rc=$?"
cut -c1-330 $out/r2c_ab_c2.jsonl
