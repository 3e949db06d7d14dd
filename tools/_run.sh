#!/bin/bash
T=${1:-s49}
mkdir -p gpurun_out
timeout 300 python bench.py --quick --steps 60 --warmup 10 --dump-ops gpurun_out/${T}_ops.json > gpurun_out/${T}_q.json 2>gpurun_out/${T}.err; cut -c1-200 gpurun_out/${T}_q.json
python tools/show_ops.py gpurun_out/${T}_ops.json | grep -E "preproc|pred|decode|sort|nms|total"
timeout 600 python bench.py --workload streams64 --steps 60 --warmup 10 > gpurun_out/${T}_streams.json 2>>gpurun_out/${T}.err; echo "rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/s49_streams.json"))
print(d["value"], d["ms_per_step"], d["per_step"], d["kernels"])
PY
tail -3 gpurun_out/${T}.err
