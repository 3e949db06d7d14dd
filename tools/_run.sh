#!/bin/bash
T=${1:-f07}
mkdir -p gpurun_out
for L in 5 4; do for K in 20 200; do
echo "lanes=$L K=$K"; TCD_LANES=$L timeout 600 python bench.py --steps $K --warmup 5 --no-aux --no-cpu-baseline > gpurun_out/${T}_l${L}_k$K.json 2>>gpurun_out/${T}.err
python tools/show_bench.py gpurun_out/${T}_l${L}_k$K.json 2>/dev/null | sed -n 1,5p | cut -c1-120
done; done
