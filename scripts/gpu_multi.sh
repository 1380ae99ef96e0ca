#!/bin/bash
# Multi-GPU session (gpurun --gpus N): the bench contract under torchrun, both arms, both splits.
# usage: bash scripts/gpu_multi.sh N [tag]
N=${1:-2}; TAG=${2:-m$N}
OUT=gpurun_out; mkdir -p $OUT
nvidia-smi -L > $OUT/${TAG}_gpu.txt
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for split in samples rows; do
  echo "== bench --gpus $N --split $split"
  timeout 900 $RUN bench.py --gpus $N --steps 3 --warmup 3 --split $split > $OUT/${TAG}_bench_$split.json 2> $OUT/${TAG}_bench_$split.err
  echo "rc=$?"; tail -1 $OUT/${TAG}_bench_$split.json; tail -3 $OUT/${TAG}_bench_$split.err
done
echo "== bench N=1 for the ratio"
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_n1.json 2> $OUT/${TAG}_bench_n1.err; tail -1 $OUT/${TAG}_bench_n1.json
echo "== reference arm under torchrun (rank 0 only works)"
timeout 600 $RUN bench.py --impl reference --gpus $N --steps 1 --warmup 0 > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench_ref.err; echo "rc=$?"; tail -1 $OUT/${TAG}_bench_ref.json
