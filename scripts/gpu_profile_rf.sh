#!/bin/bash
# On the GPU box: one full ncu capture of the fused RandomForest kernel (C3 shape, 400,000 dialogues).
TAG=${1:-rf}
mkdir -p gpurun_out
CMD="python scripts/bench_configs.py --configs c3 --c3-rows 400000"
$CMD > gpurun_out/plain_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"k_main<.int.3, .bool.1, .int.8" -s 3 -c 1 -f -o gpurun_out/$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo "profile exit $?"; tail -2 gpurun_out/ncu_full_$TAG.log; cat gpurun_out/plain_$TAG.log | tail -2
