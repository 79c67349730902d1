#!/bin/bash
# On the GPU box: full ncu captures of the featurize kernels: k_main<MODE_CSR, CLEAN, TL=8> (text -> staging) and k_gather_csr.
TAG=${1:-csr}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --rows 200000 --cpu-rows 1000 --stages"
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"k_main<.int.0, .bool.1, .int.8" -s 2 -c 1 -f -o gpurun_out/$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo "profile exit $?"; tail -2 gpurun_out/ncu_full_$TAG.log
ncu --set full --clock-control none --import-source on -k regex:k_gather_csr -s 2 -c 1 -f -o gpurun_out/${TAG}_gather $CMD > gpurun_out/ncu_full_${TAG}_gather.log 2>&1
echo "profile exit $?"; tail -2 gpurun_out/ncu_full_${TAG}_gather.log
