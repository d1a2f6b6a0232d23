#!/bin/bash
# A/B of the copy-engine X pull (PCLIP_PEER_PULL) under torchrun: usage tools/gpu_pull_ab.sh <N>
N=${1:-2}
mkdir -p gpurun_out
run() { # name, env
  timeout 170 env $2 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port $((29500 + RANDOM % 400)) bench.py --gpus $N --steps 30 --warmup 8 --no-extras \
    > gpurun_out/pull_$1_n$N.json 2> gpurun_out/pull_$1_n$N.err
  echo "rc=$? $1"; tail -c 1500 gpurun_out/pull_$1_n$N.json; tail -5 gpurun_out/pull_$1_n$N.err
}
run pull "PCLIP_PEER_PULL=1"
run nccl "PCLIP_PEER_PULL=0"
