#!/bin/bash
# On the GPU box: full bench (C2), the reference arm, then an ncu launch list of a short run.
mkdir -p gpurun_out
python bench.py --steps 10 --warmup 3 --stages > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; tail -3 gpurun_out/bench.err; cat gpurun_out/bench.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit $?"; tail -3 gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
python bench.py --steps 2 --warmup 3 --rows 200000 --cpu-rows 1000 > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --rows 200000 --cpu-rows 1000 > gpurun_out/ncu.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/ncu.log
