#!/bin/bash
# On an N-GPU box: RF training config C5 (strong: 10M rows total; weak: 10M rows per GPU).   Usage: bash scripts/gpu_train_n.sh N
N=${1:-2}
mkdir -p gpurun_out
: > gpurun_out/train_n$N.jsonl
for MODE in "" "--train-weak"; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N scripts/bench_configs.py --configs c5 --train-rows 10000000 --train-trees 100 --train-depth 10 $MODE >> gpurun_out/train_n$N.jsonl 2>> gpurun_out/train_n$N.err
done
cut -c1-330 gpurun_out/train_n$N.jsonl
