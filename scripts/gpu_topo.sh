#!/bin/bash
# On an N-GPU box: PCIe/NUMA topology, then bench.py under torchrun.   Usage: bash scripts/gpu_topo.sh N
N=${1:-8}
mkdir -p gpurun_out
{ nvidia-smi topo -m; lscpu | grep -i "numa\|socket\|^CPU(s)"; nproc; for d in /sys/bus/pci/devices/*; do if [ "$(cat $d/class 2>/dev/null)" = "0x030200" ]; then echo "$d numa=$(cat $d/numa_node)"; fi; done; free -g | head -2; } > gpurun_out/topo_n$N.txt 2>&1
cat gpurun_out/topo_n$N.txt | head -40
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench exit $?"; tail -3 gpurun_out/bench_n$N.err; tail -c 700 gpurun_out/bench_n$N.json
