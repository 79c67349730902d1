#!/bin/bash
# On an 8-GPU box: scaling of training (strong: 10M rows total; weak: 10M rows per GPU) at N = 1, 2, 4, 8.
mkdir -p gpurun_out
: > gpurun_out/train_scale.jsonl
for N in 1 2 4 8; do
  for MODE in "" "--train-weak"; do
    if [ $N -eq 1 ]; then
      python scripts/bench_configs.py --configs c5 --train-rows 10000000 --train-trees 100 --train-depth 10 $MODE >> gpurun_out/train_scale.jsonl 2>> gpurun_out/train_scale.err
    else
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N scripts/bench_configs.py --configs c5 --train-rows 10000000 --train-trees 100 --train-depth 10 $MODE >> gpurun_out/train_scale.jsonl 2>> gpurun_out/train_scale.err
    fi
  done
done
cat gpurun_out/train_scale.jsonl
