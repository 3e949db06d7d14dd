#!/bin/bash
T=${1:-s50}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_stream_bank.py -q -x > gpurun_out/${T}_bank.log 2>&1; echo "bank rc=$?"; tail -3 gpurun_out/${T}_bank.log
timeout 600 python bench.py --workload streams64 --steps 60 --warmup 10 > gpurun_out/${T}_streams.json 2>gpurun_out/${T}.err; echo "rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/s50_streams.json"))
print(d["value"], d["ms_per_step"], d["per_step"], d["kernels"])
PY
tail -3 gpurun_out/${T}.err
