#!/bin/bash
T=${1:-s52}
mkdir -p gpurun_out
for i in 1 2 3; do timeout 300 python bench.py --quick --steps 20 --warmup 5 2>>gpurun_out/${T}.err | cut -c1-110; done
timeout 900 python bench.py --steps 20 --warmup 5 --no-aux > gpurun_out/${T}_bench20.json 2>> gpurun_out/${T}.err; python tools/show_bench.py gpurun_out/${T}_bench20.json | head -6
