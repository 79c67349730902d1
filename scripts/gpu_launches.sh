#!/bin/bash
# On the GPU box: ncu launch list (per-launch durations) of a short bench run with the stage timings.
TAG=${1:-l}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --rows 200000 --cpu-rows 1000 --stages"
$CMD > gpurun_out/plain_$TAG.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:k_ -c 300 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_$TAG.log 2>&1
echo "exit $?"
