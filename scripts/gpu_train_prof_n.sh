#!/bin/bash
# On an N-GPU box: RF training config C5 (strong: 10M rows total) with the trainer's phase timers.   Usage: bash scripts/gpu_train_prof_n.sh N
N=${1:-2}
mkdir -p gpurun_out
: > gpurun_out/trainprof_n$N.jsonl
for EXTRA in "" "--train-profile"; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2954$N scripts/bench_configs.py --configs c5 --train-rows 10000000 --train-trees 100 --train-depth 10 $EXTRA >> gpurun_out/trainprof_n$N.jsonl 2>> gpurun_out/trainprof_n$N.err
done
cat gpurun_out/trainprof_n$N.jsonl
