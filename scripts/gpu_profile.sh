#!/bin/bash
# On the GPU box: ncu launch list of this repo's kernels + one full capture of the dominant kernel.
# Usage: bash scripts/gpu_profile.sh [tag]
TAG=${1:-prof}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --rows 200000 --cpu-rows 1000 --stages"
$CMD > gpurun_out/plain_$TAG.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_ -c 200 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_main -s 3 -c 1 -f -o gpurun_out/$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo "profile exit $?"; tail -2 gpurun_out/ncu_full_$TAG.log
