#!/bin/bash
# On an N-GPU box: bench.py and the training config under torchrun.   Usage: bash scripts/gpu_multi.sh N
N=${1:-2}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench exit $?"; tail -3 gpurun_out/bench_n$N.err; cat gpurun_out/bench_n$N.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 scripts/bench_configs.py --configs c3,c5 --train-rows 10000000 --train-trees 100 --train-depth 10 > gpurun_out/configs_n$N.jsonl 2> gpurun_out/configs_n$N.err; echo "configs exit $?"; tail -3 gpurun_out/configs_n$N.err; cat gpurun_out/configs_n$N.jsonl
