#!/bin/bash
# 8-GPU box: C4 and C2 at N=8 and N=4 (torchrun, one rank per GPU), the C caller on 8 devices, the multi-device test
set -u
out=gpurun_out
run() { # n workload extra...
  n=$1; wlname=$2; shift 2
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n --workload $wlname "$@" > $out/r2_bench_${wlname}_n$n.json 2> $out/r2_bench_${wlname}_n$n.err
  echo "$wlname n=$n rc=$?"; tail -c 1500 $out/r2_bench_${wlname}_n$n.json | head -c 1500; echo
}
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8 > $out/r2m_gpus.txt
run 8 c4 --steps 2 --warmup 3 --no-cpu-baseline
run 4 c4 --steps 2 --warmup 3 --no-cpu-baseline
run 8 c2 --steps 5 --warmup 3 --no-cpu-baseline
run 4 c2 --steps 5 --warmup 3 --no-cpu-baseline
timeout 300 tools/bin/terra_headless --scene field --width 1920 --height 1080 --spp 256 --passes 3 --devices 0,1,2,3,4,5,6,7 --out $out/r2m_field_8gpu.ppm > $out/r2m_headless_8.log 2>&1; echo "headless 8 rc=$?"; cat $out/r2m_headless_8.log
timeout 300 python -m pytest tests/test_gpu_state.py -x -q -m gpu > $out/r2m_pytest_state.log 2>&1; echo "pytest rc=$?"; tail -3 $out/r2m_pytest_state.log
rm -f $out/r2m_field_8gpu.ppm
