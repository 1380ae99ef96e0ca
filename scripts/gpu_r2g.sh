#!/bin/bash
# two GPUs: the in-library multi-device path (test + headless caller) and the torchrun bench at N=2
set -u
out=gpurun_out
nvidia-smi -L > $out/r2g_gpu.txt
python -m pytest tests/test_gpu_state.py -m gpu -q -x > $out/r2g_pytest_state.log 2>&1; echo "pytest state rc=$?"
tail -5 $out/r2g_pytest_state.log
./tools/bin/terra_headless --scene field --width 1920 --height 1080 --spp 256 --bounces 4 --passes 3 --out $out/r2g_field_1gpu.ppm > $out/r2g_headless_1.log 2>&1; echo "headless 1 rc=$?"
cat $out/r2g_headless_1.log
./tools/bin/terra_headless --scene field --width 1920 --height 1080 --spp 256 --bounces 4 --passes 3 --devices 0,1 --out $out/r2g_field_2gpu.ppm > $out/r2g_headless_2.log 2>&1; echo "headless 2 rc=$?"
cat $out/r2g_headless_2.log
rm -f $out/r2g_field_*.ppm
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 4 --warmup 3 > $out/r2g_bench_n2.json 2> $out/r2g_bench_n2.err; echo "bench n2 rc=$?"
tail -c 900 $out/r2g_bench_n2.json; tail -3 $out/r2g_bench_n2.err
python bench.py --gpus 1 --steps 4 --warmup 3 --no-cpu-baseline > $out/r2g_bench_n1.json 2> $out/r2g_bench_n1.err; echo "bench n1 rc=$?"
python - <<'PY'
import json
for n in (1,2):
    try:
        d=json.loads(open(f"gpurun_out/r2g_bench_n{n}.json").read().strip().splitlines()[-1])
        print(n, "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms", round(d["ms_per_step"],2), "e2e ms", round(d["e2e"]["ms_per_step"],2))
    except Exception as e: print(n, "ERR", e)
PY
