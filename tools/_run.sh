#!/bin/bash
# usage: tools/_run.sh TAG  -- GPU box: tests, host probe
T=${1:-s31}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|FAILED|Error|confident oracle" gpurun_out/${T}_pytest.log | head -40
timeout 300 python tools/host_probe2.py > gpurun_out/${T}_probe.log 2>&1; echo "probe rc=$?"
head -c 4500 gpurun_out/${T}_probe.log
