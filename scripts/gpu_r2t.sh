#!/bin/bash
set -u
out=gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_lbvh.py tests/test_gpu_state.py -x -q -m gpu > $out/r2t_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 $out/r2t_pytest.log
