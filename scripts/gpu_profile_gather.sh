#!/bin/bash
TAG=${1:-gather}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --rows 200000 --cpu-rows 1000 --stages"
ncu --set full --clock-control none --import-source on -k regex:k_gather_csr -s 2 -c 1 -f -o gpurun_out/${TAG} $CMD > gpurun_out/ncu_full_${TAG}.log 2>&1
echo "profile exit $?"; tail -2 gpurun_out/ncu_full_${TAG}.log
