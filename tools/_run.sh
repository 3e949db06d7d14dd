#!/bin/bash
T=${1:-f02}
mkdir -p gpurun_out
for i in 1 2; do
timeout 600 python bench.py --no-aux --no-cpu-baseline > gpurun_out/${T}_b$i.json 2> gpurun_out/${T}_b$i.err; echo "rc=$?"
python tools/show_bench.py gpurun_out/${T}_b$i.json 2>/dev/null | sed -n 1,6p
done
nproc; lscpu | grep -E "Model name|MHz" | head -3
