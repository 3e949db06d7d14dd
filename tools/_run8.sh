#!/bin/bash
# 8-GPU box: both workloads under torchrun
T=${1:-s47}
mkdir -p gpurun_out
for N in 8; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 40 --warmup 5 --no-aux > gpurun_out/${T}_bench_n$N.json 2> gpurun_out/${T}_bench_n$N.err; echo "bench N=$N rc=$?"
  tail -c 400 gpurun_out/${T}_bench_n$N.err
  python tools/show_bench.py gpurun_out/${T}_bench_n$N.json
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --workload streams64 --steps 40 --warmup 10 > gpurun_out/${T}_streams_n$N.json 2> gpurun_out/${T}_streams_n$N.err; echo "streams N=$N rc=$?"
  tail -c 300 gpurun_out/${T}_streams_n$N.err
  cut -c1-400 gpurun_out/${T}_streams_n$N.json
done
nproc; lscpu | grep -E "NUMA|Model name|Socket" | head -8; nvidia-smi topo -m | head -12
