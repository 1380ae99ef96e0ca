#!/bin/bash
set -u
out=gpurun_out
python scripts/diag_textured.py > $out/r2b_diag_textured.log 2>&1; echo "diag rc=$?"
cat $out/r2b_diag_textured.log
python -m pytest tests -m gpu -q -s --deselect "tests/test_gpu_baseline_scenes.py::test_baseline_scene_image_matches_cpu_renderer" > $out/r2b_pytest_gpu.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|FAILED|ERROR" $out/r2b_pytest_gpu.log | tail -30
python -m pytest tests/test_gpu_baseline_scenes.py -m gpu -q -s -k "textured" > $out/r2b_pytest_textured.log 2>&1; echo "pytest textured rc=$?"
grep -E "passed|failed|FAILED|ERROR|rmse|mean rel" $out/r2b_pytest_textured.log | tail -30
