#!/bin/bash
T=${1:-s40}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_stream_bank.py tests/test_gpu_motion.py -q -x > gpurun_out/${T}_bank.log 2>&1; echo "bank rc=$?"
grep -E "passed|failed|FAILED|Error|assert" gpurun_out/${T}_bank.log | head -20
timeout 600 python bench.py --quick --steps 60 --warmup 10 --dump-ops gpurun_out/${T}_ops.json > gpurun_out/${T}_q.json 2>gpurun_out/${T}.err; cut -c1-200 gpurun_out/${T}_q.json
python tools/show_ops.py gpurun_out/${T}_ops.json | grep -E "pred|decode|total"
