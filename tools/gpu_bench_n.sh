#!/bin/bash
# One headline bench run under torchrun with a hard time limit: tools/gpu_bench_n.sh <N> [extra bench args]
N=${1:-2}; shift
mkdir -p gpurun_out
timeout 170 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
  --master-port $((29500 + RANDOM % 400)) bench.py --gpus $N --steps 30 --warmup 8 "$@" \
  > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "rc=$?"; tail -c 2500 gpurun_out/bench_n$N.json; tail -3 gpurun_out/bench_n$N.err
