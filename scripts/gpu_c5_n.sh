#!/bin/bash
# On an N-GPU box: config C5 (strong + weak, with phase split) through scripts/bench_configs.py.   Usage: bash scripts/gpu_c5_n.sh N [rows]
N=${1:-2}
ROWS=${2:-10000000}
mkdir -p gpurun_out
if [ "$N" = "1" ]; then
  python scripts/bench_configs.py --configs c5 --train-rows $ROWS > gpurun_out/c5_n$N.jsonl 2> gpurun_out/c5_n$N.err
else
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2955$N scripts/bench_configs.py --configs c5 --train-rows $ROWS > gpurun_out/c5_n$N.jsonl 2> gpurun_out/c5_n$N.err
fi
echo "exit $?"; tail -3 gpurun_out/c5_n$N.err; cat gpurun_out/c5_n$N.jsonl
