#!/bin/bash
T=${1:-f04}
N=${2:-8}
mkdir -p gpurun_out
for i in 1 2 3 4; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$i bench.py --gpus $N --steps 20 --warmup 5 --quick 2> gpurun_out/${T}_q$i.err | cut -c1-240
done
