#!/bin/bash
# On the GPU box: RF training config C5 on one GPU with the trainer's phase timers (synchronising; slower than the plain run).
mkdir -p gpurun_out
python scripts/bench_configs.py --configs c5 --train-rows 10000000 --train-trees 100 --train-depth 10 > gpurun_out/c5_plain.jsonl 2> gpurun_out/c5_plain.err; echo "exit $?"; cat gpurun_out/c5_plain.jsonl
python scripts/bench_configs.py --configs c5 --train-rows 10000000 --train-trees 100 --train-depth 10 --train-profile > gpurun_out/c5_profile.jsonl 2> gpurun_out/c5_profile.err; echo "exit $?"; cat gpurun_out/c5_profile.jsonl
