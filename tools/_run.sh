#!/bin/bash
T=${1:-s68}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_ingest.py tests/test_gpu_stream_bank.py -q -x > gpurun_out/${T}_py.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/${T}_py.log
timeout 300 python - <<'PY' 2>&1 | tail -5
import sys, json
sys.argv=["bench.py"]
import bench
print(json.dumps(bench.ingest_side(0)))
PY
