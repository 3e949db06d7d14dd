#!/bin/bash
T=${1:-s67}
N=${2:-8}
mkdir -p gpurun_out
TCD_BENCH_AB=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 --quick > gpurun_out/${T}_ab_n${N}.json 2> gpurun_out/${T}_ab_n${N}.err; echo "rc=$?"
grep "^AB" gpurun_out/${T}_ab_n${N}.err
cut -c1-300 gpurun_out/${T}_ab_n${N}.json
