#!/bin/bash
T=${1:-s65}
mkdir -p gpurun_out
for dbg in 0 1 3; do
echo "dbg=$dbg"
TCD_CONV_DBG=$dbg TCD_LANES=1 timeout 300 python bench.py --quick --batch 32 --obj-bias=-8 --steps 5 --warmup 2 --dump-ops gpurun_out/${T}_ops.json 2>>gpurun_out/${T}.err | cut -c1-100
python tools/show_ops.py gpurun_out/${T}_ops.json | grep -E "pred|stem|dark2.0|dark3.1.conv12|dark4.0|total"
done
