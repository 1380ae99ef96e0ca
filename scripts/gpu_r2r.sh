#!/bin/bash
# k_trace_v7 (4-wide tree): parity first, then A/B against v6
set -u
out=gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_lbvh.py tests/test_gpu_state.py tests/test_zz_no_alloc.py -x -q -m gpu > $out/r2r_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 $out/r2r_pytest.log
timeout 600 python scripts/ab_trace.py c3 > $out/r2r_ab_c3.jsonl 2> $out/r2r_ab_c3.err; echo "c3 rc=$?"; cat $out/r2r_ab_c3.jsonl; tail -3 $out/r2r_ab_c3.err
timeout 600 python scripts/ab_trace.py c2 --steps 5 > $out/r2r_ab_c2.jsonl 2> $out/r2r_ab_c2.err; echo "c2 rc=$?"; cat $out/r2r_ab_c2.jsonl; tail -3 $out/r2r_ab_c2.err
timeout 900 python scripts/ab_trace.py c3 --tris 10000000 --only "v" > $out/r2r_ab_c3_10m.jsonl 2> $out/r2r_ab_c3_10m.err; echo "c3 10m rc=$?"; head -3 $out/r2r_ab_c3_10m.jsonl; tail -3 $out/r2r_ab_c3_10m.err
