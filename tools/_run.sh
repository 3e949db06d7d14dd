#!/bin/bash
T=${1:-s69}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_enhance.py tests/test_ingest.py -q -x > gpurun_out/${T}_py.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/${T}_py.log
