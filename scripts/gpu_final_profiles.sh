#!/bin/bash
# On the GPU box: the round's profile set -- launch list, full captures of the headline kernel (k_main<DT>), the featurize
# kernels (k_main<CSR>, k_gather_copy) and the CSR scorer.   Usage: bash scripts/gpu_final_profiles.sh TAG
TAG=${1:-r02}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --rows 200000 --cpu-rows 1000 --configs ''"
$CMD > gpurun_out/plain_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_$TAG.log; exit 1; }
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:k_ -c 300 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/ncu_$TAG.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"k_main<.int.2" -s 3 -c 1 -f -o gpurun_out/${TAG}_k_main_dt $CMD > gpurun_out/ncu_full_dt_$TAG.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"k_main<.int.0, .bool.1, .int.8" -s 2 -c 1 -f -o gpurun_out/${TAG}_k_main_csr $CMD > gpurun_out/ncu_full_csr_$TAG.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_gather_copy -s 2 -c 1 -f -o gpurun_out/${TAG}_k_gather_copy $CMD > gpurun_out/ncu_full_gather_$TAG.log 2>&1
echo "profiles exit $?"; ls -la gpurun_out/${TAG}_*
