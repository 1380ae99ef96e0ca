#!/bin/bash
# Multi-GPU check of the final kernels on an N-GPU box: the multi-device tests, the torchrun bench line(s) the driver's
# SCALE run launches, and the C caller through ONE terra_render call.   bash scripts/gpu_multi.sh TAG N [workload ...]
set -u
TAG=$1; N=$2; shift 2
out=gpurun_out; mkdir -p $out
run() { n=$1; wlname=$2; shift 2
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n --workload $wlname "$@" > $out/${TAG}_bench_${wlname}_n$n.json 2> $out/${TAG}_bench_${wlname}_n$n.err
  echo "$wlname n=$n rc=$?"; python - <<PY
import json
j=json.loads(open("$out/${TAG}_bench_${wlname}_n$n.json").read().strip().splitlines()[-1])
print("value", round(j["value"]), "e2e", round(j["e2e"]["value"]), "ms", round(j["ms_per_step"],2), "e2e ms", round(j["e2e"].get("ms_per_step", 0),2))
PY
}
timeout 600 python -m pytest tests/test_gpu_state.py tests/test_gpu_render.py -x -q -m gpu -k "multi or split or device" > $out/${TAG}_pytest_multi.log 2>&1; echo "pytest rc=$?"; tail -3 $out/${TAG}_pytest_multi.log
for wl in "${@:-c2}"; do
  steps=6; [ "$wl" = c4 ] && steps=2
  run $N $wl --steps $steps --warmup 3 --no-cpu-baseline
done
devs=$(seq -s, 0 $((N - 1)))
timeout 300 tools/bin/terra_headless --scene field --width 1920 --height 1080 --spp 256 --passes 4 --devices $devs --out $out/${TAG}_field.ppm > $out/${TAG}_headless_$N.log 2>&1; echo "headless $N rc=$?"; cat $out/${TAG}_headless_$N.log
rm -f $out/${TAG}_field.ppm
