#!/bin/bash
set -u
out=gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu > $out/r2s_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $out/r2s_pytest.log
for t in 1000000 3000000 10000000; do
timeout 900 python scripts/ab_trace.py c3 --tris $t --only "tree" >> $out/r2s_ab_c3.jsonl 2>> $out/r2s_ab_c3.err; echo "c3 $t rc=$?"
done
cat $out/r2s_ab_c3.jsonl; tail -3 $out/r2s_ab_c3.err
timeout 600 python scripts/ab_trace.py c2 --steps 5 > $out/r2s_ab_c2.jsonl 2> $out/r2s_ab_c2.err; echo "c2 rc=$?"; cat $out/r2s_ab_c2.jsonl
