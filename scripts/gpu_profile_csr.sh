#!/bin/bash
# On the GPU box: one full ncu capture of the featurize (text -> CSR staging) kernel k_rows<MODE_CSR, CLEAN, !BIG, TL=8>.
TAG=${1:-csr}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --rows 200000 --cpu-rows 1000 --stages"
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"k_rows<.int.0, .bool.1, .int.0, .int.8>" -s 2 -c 1 -f -o gpurun_out/$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo "profile exit $?"; tail -2 gpurun_out/ncu_full_$TAG.log
