#!/bin/bash
set -u
out=gpurun_out
n=${1:-2}
timeout 300 python -m pytest tests/test_gpu_state.py -x -q -m gpu > $out/r2o_pytest_state_$n.log 2>&1; echo "pytest rc=$?"; tail -15 $out/r2o_pytest_state_$n.log
for chunk in 16777216 1000000000000; do
TERRA_B200_FANOUT_CHUNK=$chunk timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 6 --warmup 3 --no-cpu-baseline > $out/r2o_bench_c2_n${n}_chunk$chunk.json 2> $out/r2o_bench_c2_n${n}.err
echo "chunk $chunk n=$n rc=$?"; python - <<PY
import json
j=json.loads(open("$out/r2o_bench_c2_n${n}_chunk$chunk.json").read().strip().splitlines()[-1])
print("value", round(j["value"]), "e2e", round(j["e2e"]["value"]), "ms", round(j["ms_per_step"],2), "e2e ms", round(j["e2e"]["ms_per_step"],2))
PY
done
