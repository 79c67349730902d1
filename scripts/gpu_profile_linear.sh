#!/bin/bash
# On the GPU box: full ncu capture of the fused LogisticRegression kernel (config C1's device-resident pass, 200k dialogues).
mkdir -p gpurun_out
CMD="python scripts/bench_configs.py --configs c1 --c1-rows 200000"
$CMD > gpurun_out/plain_linear.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_linear.log; exit 1; }
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"k_main<.int.1, .bool.1, .int.8" -s 2 -c 1 -f -o gpurun_out/r02_k_main_lr $CMD > gpurun_out/ncu_full_lr.log 2>&1
echo "exit $?"; tail -2 gpurun_out/ncu_full_lr.log | cut -c1-200
