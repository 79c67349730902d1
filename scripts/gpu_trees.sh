#!/bin/bash
# On the GPU box: tree-model parity tests, then the RF (C3) and GBT (C4) configs on one GPU.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "tree or golden or batcher or no_alloc" > gpurun_out/pytest_trees.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_trees.log
python scripts/bench_configs.py --configs c3,c4 > gpurun_out/configs_trees.jsonl 2> gpurun_out/configs_trees.err; echo "configs exit $?"; tail -3 gpurun_out/configs_trees.err; cat gpurun_out/configs_trees.jsonl
