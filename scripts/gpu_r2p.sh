#!/bin/bash
set -u
out=gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > $out/r2p_pytest_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -15 $out/r2p_pytest_noalloc_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu > $out/r2p_pytest_gpu.log 2>&1; echo "suite rc=$?"; tail -8 $out/r2p_pytest_gpu.log
