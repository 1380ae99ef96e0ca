#!/bin/bash
# 8-GPU box again: the fan-out issued by one host thread per device. C2 at N=8 / N=4, the C caller, the multi-device test.
set -u
out=gpurun_out
run() { n=$1; wlname=$2; shift 2
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n --workload $wlname "$@" > $out/r2n_bench_${wlname}_n$n.json 2> $out/r2n_bench_${wlname}_n$n.err
  echo "$wlname n=$n rc=$?"; python - <<PY
import json
j=json.loads(open("$out/r2n_bench_${wlname}_n$n.json").read().strip().splitlines()[-1])
print("value", round(j["value"]), "e2e", round(j["e2e"]["value"]), "ms", round(j["ms_per_step"],2), "e2e ms", round(j["e2e"]["ms_per_step"],2))
PY
}
timeout 300 python -m pytest tests/test_gpu_state.py -x -q -m gpu > $out/r2n_pytest_state.log 2>&1; echo "pytest rc=$?"; tail -3 $out/r2n_pytest_state.log
run 8 c2 --steps 6 --warmup 3 --no-cpu-baseline
run 4 c2 --steps 6 --warmup 3 --no-cpu-baseline
run 2 c2 --steps 6 --warmup 3 --no-cpu-baseline
timeout 300 tools/bin/terra_headless --scene field --width 1920 --height 1080 --spp 256 --passes 4 --devices 0,1,2,3,4,5,6,7 --out $out/r2n_field_8gpu.ppm > $out/r2n_headless_8.log 2>&1; echo "headless 8 rc=$?"; cat $out/r2n_headless_8.log
rm -f $out/r2n_field_8gpu.ppm
run 2 c4 --steps 2 --warmup 3 --no-cpu-baseline
