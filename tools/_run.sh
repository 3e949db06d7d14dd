#!/bin/bash
T=${1:-s71}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_jpeg.py -q -x > gpurun_out/${T}_py.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/${T}_py.log
