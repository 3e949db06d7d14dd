#!/bin/bash
T=${1:-s46}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|FAILED|Error|max err" gpurun_out/${T}_pytest.log | head -20
timeout 300 python bench.py --quick --steps 120 --warmup 10 --dump-ops gpurun_out/${T}_ops.json > gpurun_out/${T}_q.json 2>gpurun_out/${T}.err; cut -c1-230 gpurun_out/${T}_q.json
TCD_LANES=1 TCD_ENGINES=1 timeout 300 python bench.py --quick --batch 32 --steps 10 --warmup 3 --dump-ops gpurun_out/${T}_ops32.json > gpurun_out/${T}_q32.json 2>>gpurun_out/${T}.err; cut -c1-230 gpurun_out/${T}_q32.json
python tools/show_ops.py gpurun_out/${T}_ops32.json | awk '$2>35 || /total/'
